// sm_100a kernels of the 1090 MHz path.
//
//   scan_kernel      fused: uint8 I/Q -> magnitudes (registers -> shared tile with a
//                    290-sample lookahead halo) -> preamble detector over every start
//                    position (warp ballots build rising/falling-edge bit masks, the five
//                    preamble patterns are evaluated 32 positions at a time) -> five-phase
//                    bit slicer + CRC-24 syndromes + syndrome-table repair -> compaction
//                    of (position, phase) records that can still reach the accept
//                    threshold.  Replaces magnitude_uc8.c:30-63 + demod_2400.c:134-347 +
//                    mode_s.c:252-466 (state-independent part).
//                    Each CTA owns a contiguous run of tiles (one "segment") and writes its records
//                    in (position, phase) order, so the chunk comes out in stream order.
//   filter_kernel    per segment, in place and order preserving: drops Address/Parity records
//                    whose address can never be in the ICAO filter (bitmap built by the scan).
//   evaluate_kernel  one warp per surviving record: re-slices it from the I/Q samples and
//                    writes the fully evaluated record, in stream order, straight into mapped
//                    host memory for the ordered host pass (demod_2400.c:349-423 is
//                    sequential by definition).
//   carry_kernel     keeps the last 326 magnitudes for the next feed (fifo.c:187-196).
//   magnitude_kernel stage-level I/Q -> magnitude (starch_magnitude_uc8).
#include "adsbcu_device.cuh"

namespace adsbcu {

namespace {

#ifdef ADSBCU_PHASE_TIMING           // debug: cycles between group barriers, summed over all groups (thread 0 of each)
__device__ unsigned long long g_phase_cycles[8];
// PHASE_MARK(k): the interval that ends here belongs to slot k
#define PHASE_MARK(k) do { if (tid == 0) { const long long t_now = clock64(); if (t_prev) atomicAdd(&g_phase_cycles[k], (unsigned long long)(t_now - t_prev)); t_prev = t_now; } } while (0)
#else
#define PHASE_MARK(k) do { } while (0)
#endif
constexpr int kPreCap = kTileSamples / 4;      // positions passing the edge masks, per tile
constexpr int kCandCap = kTileSamples / 8;     // preambles kept per tile (overflow -> error, never silently dropped)
constexpr int kCandChunk = kGroupThreads;      // preambles per slicer round
constexpr int kQueueCap = 5 * kCandChunk;      // (preamble, phase) items per round
constexpr int kKeepCap = kGroups > 4 ? kTileSamples / 10 : kTileSamples / 8;     // kept (position, phase) records per tile
constexpr int kBuckets = kTileSamples / 48, kBucketShift = 9; // phase D counting sort: 64 positions x 8 phase slots per bucket
constexpr int kPosWords = (kTile + 31) / 32;   // 32-position words per tile
constexpr int kPlaneStride = kGroupThreads + 16;   // symbol-plane fields per plane (+ zeroed slack for the 128-bit reads)
static_assert(kPosWords + 2 <= kGroupThreads, "phase B: one thread per 32 start positions");

// Scratch that is live in different phases of a tile shares storage.
union GroupScratch {
    struct {                                   // phases A..B2
        __align__(16) uint8_t lbits[kTileSamples / 8 + 16];   // bit i: mag[i] < mag[i+1]
        __align__(16) uint8_t gbits[kTileSamples / 8 + 16];   // bit i: mag[i] > mag[i+1]
        uint16_t pre[kPreCap];
    } edge;
    uint2 sorted[kKeepCap];                        // phase D
};
struct GroupSmem {
    __align__(16) uint16_t mag[kTileSamples + kTilePad];
    // Symbol planes.  The slicer decision at "time" t = 5 * sample + sub (sub = correlator phase 0..4) only
    // depends on t, and symbol k of a message starting at time t0 sits at t0 + 12 k: a message is therefore
    // 56 / 112 CONSECUTIVE bits of the plane t0 mod 12.  plane[c][v] holds, MSB first in its low 10 bits, the
    // 10 decisions of plane c inside the 24 samples (120 fifths) of thread v.
    uint16_t plane[12][kPlaneStride];
    GroupScratch u;
    union {
        uint16_t cand[kCandCap];               // phases B2..C
        uint32_t hist[3 * kBuckets];           // phase D
    };
    uint32_t queue[kQueueCap];                 // 112-bit items from the front, 56-bit items from the back: pos | (phase-4) << 16
    uint2    keep[kKeepCap];                   // x: pos << 3 | phase-4 (sort key), y: PhaseRec.info
    uint32_t npre, ncand, nq[2], nkeep;
};
struct ScanSmem {
    __align__(16) uint16_t lut[32768];         // DevTables::mag_lut
    uint32_t  crcb[11 * 256];
    float     sq[256];                         // DevTables::sq, for samples outside the central LUT
    uint32_t  occ_long[kSyndSlotsLong / 32], occ_short[kSyndSlotsShort / 32];   // DevTables::occ_*
    GroupSmem grp[kGroups];
};
static_assert(sizeof(ScanSmem) <= 227 * 1024, "scan kernel shared memory exceeds the sm_100 per-CTA limit");

__device__ __forceinline__ uint32_t funnel_r(uint32_t lo, uint32_t hi, int sh) { return __funnelshift_r(lo, hi, sh); }
__device__ __forceinline__ void group_sync(int g)
{
#ifdef ADSBCU_EMU
    emu::barrier(g + 1, kGroupThreads);
#else
    asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "n"(kGroupThreads) : "memory");
#endif
}
// packed magnitudes {m(I0,Q0), m(I1,Q1)} of the two samples in a 32-bit word {I0,Q0,I1,Q1}; both must be
// inside the central LUT (see lut_offset)
__device__ __forceinline__ uint32_t lut_pair(const uint16_t *lut, uint32_t w)
{
    const uint32_t u = (w ^ ((w >> 7) & 0x003E003Eu)) & 0x7F7F7F7Fu;   // both element indices at once
    return uint32_t(lut[u & 0xFFFFu]) | (uint32_t(lut[u >> 16]) << 16);
}
// every byte of w in [64, 192)  <=>  bit7 != bit6 in every byte
__device__ __forceinline__ uint32_t central_bits(uint32_t w) { return w ^ (w << 1); }
__device__ __forceinline__ uint32_t mag_pair_any(const float *sq, uint32_t w)
{
    return mag_iq(sq, w & 0xFFu, (w >> 8) & 0xFFu) | (mag_iq(sq, (w >> 16) & 0xFFu, w >> 24) << 16);
}
__host__ __device__ constexpr uint32_t w2(int lo, int hi) { return uint32_t(uint8_t(lo)) | (uint32_t(uint8_t(hi)) << 8); }

// Reference preamble test at p[0..18] (demod_2400.c:155-218).  `p` points into shared memory.
__device__ __forceinline__ bool preamble_ok(const uint16_t *p)
{
    const int p1 = p[1], p2 = p[2], p3 = p[3], p4 = p[4], p5 = p[5], p6 = p[6], p7 = p[7], p8 = p[8], p9 = p[9],
              p10 = p[10], p11 = p[11], p12 = p[12];
    int high;
    unsigned sig, noise;
    const bool up12 = p11 < p12;
    if (p1 > p2 && p2 < p3 && p3 > p4 && p8 < p9 && p9 > p10 && p10 < p11) {
        high = (p1 + p3 + p9 + p11 + p12) / 4; sig = p1 + p3 + p9; noise = p5 + p6 + p7;
    } else if (p1 > p2 && p2 < p3 && p3 > p4 && p8 < p9 && p9 > p10 && up12) {
        high = (p1 + p3 + p9 + p12) / 4; sig = p1 + p3 + p9 + p12; noise = p5 + p6 + p7 + p8;
    } else if (p1 > p2 && p2 < p3 && p4 > p5 && p8 < p9 && p10 > p11 && up12) {
        high = (p1 + p3 + p4 + p9 + p10 + p12) / 4; sig = p1 + p12; noise = p6 + p7;
    } else if (p1 > p2 && p3 < p4 && p4 > p5 && p9 < p10 && p10 > p11 && up12) {
        high = (p1 + p4 + p10 + p12) / 4; sig = p1 + p4 + p10 + p12; noise = p5 + p6 + p7 + p8;
    } else if (p2 > p3 && p3 < p4 && p4 > p5 && p9 < p10 && p10 > p11 && up12) {
        high = (p1 + p2 + p4 + p10 + p12) / 4; sig = p4 + p10 + p12; noise = p6 + p7 + p8;
    } else {
        return false;
    }
    if (sig * 2 < 3 * noise) return false;
    if (p5 >= high || p6 >= high || p7 >= high || p8 >= high) return false;
    if (p[14] >= high || p[15] >= high || p[16] >= high || p[17] >= high || p[18] >= high) return false;
    return true;
}

// kMagIn: the input is uint16 magnitudes (ScanArgs::mag_in, SC16 front end) instead of uint8 I/Q
template <bool kMagIn>
__global__ void __launch_bounds__(kGroups * kGroupThreads, 1) scan_kernel(ScanArgs a)
{
    ADSBCU_DYN_SMEM(smem_raw);
    ScanSmem &sm = *reinterpret_cast<ScanSmem *>(smem_raw);

    // ---- CTA prologue: magnitude LUT and CRC byte tables into shared memory (TMA bulk copies) ------
#ifdef ADSBCU_EMU
    for (uint32_t i = threadIdx.x; i < 32768u; i += blockDim.x) sm.lut[i] = a.tab->mag_lut[i];
    for (uint32_t i = threadIdx.x; i < 11u * 256u; i += blockDim.x) sm.crcb[i] = (&a.tab->crc_byte[0][0])[i];
    for (uint32_t i = threadIdx.x; i < 256u; i += blockDim.x) sm.sq[i] = a.tab->sq[i];
    for (uint32_t i = threadIdx.x; i < kSyndSlotsLong / 32; i += blockDim.x) sm.occ_long[i] = a.tab->occ_long[i];
    for (uint32_t i = threadIdx.x; i < kSyndSlotsShort / 32; i += blockDim.x) sm.occ_short[i] = a.tab->occ_short[i];
    __syncthreads();
#else
    {
        __shared__ __align__(8) unsigned long long s_bar;
        const uint32_t bar = uint32_t(__cvta_generic_to_shared(&s_bar));
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            constexpr uint32_t kLutBytes = 32768 * 2, kCrcBytes = 11 * 256 * 4, kSqBytes = 256 * 4 + kSyndSlotsLong / 8 + kSyndSlotsShort / 8;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(kLutBytes + kCrcBytes + kSqBytes) : "memory");
            for (uint32_t o = 0; o < kLutBytes; o += 32768u)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(uint32_t(__cvta_generic_to_shared(sm.lut)) + o), "l"(reinterpret_cast<const unsigned char *>(a.tab->mag_lut) + o), "r"(32768u), "r"(bar) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(uint32_t(__cvta_generic_to_shared(sm.crcb))), "l"(&a.tab->crc_byte[0][0]), "r"(kCrcBytes), "r"(bar) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(uint32_t(__cvta_generic_to_shared(sm.sq))), "l"(a.tab->sq), "r"(256u * 4u), "r"(bar) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(uint32_t(__cvta_generic_to_shared(sm.occ_long))), "l"(a.tab->occ_long), "r"(uint32_t(kSyndSlotsLong / 8 + kSyndSlotsShort / 8)), "r"(bar) : "memory");
        }
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(bar) : "memory");
    }
#endif

    const int grp = threadIdx.x / kGroupThreads, tid = threadIdx.x % kGroupThreads, lane = tid & 31;
    GroupSmem &gs = sm.grp[grp];
    uint16_t *s_mag = gs.mag;
    const uint16_t *lut = sm.lut;
    const float *s_sq = sm.sq;
    const uint32_t *s_crcb = sm.crcb;
    const uint32_t valid_long = a.tab->valid_long, valid_short = a.tab->valid_short, fixable_short = a.tab->fixable_short;
    for (int i = tid; i < 12 * (kPlaneStride - kGroupThreads); i += kGroupThreads)     // slack fields stay zero
        gs.plane[i / (kPlaneStride - kGroupThreads)][kGroupThreads + i % (kPlaneStride - kGroupThreads)] = 0;

    const int64_t n = a.nsamples;
    const uint32_t ntiles = (a.pos_end - a.pos_begin + kTile - 1) / kTile;
    // this group's segment: a contiguous run of the chunk's tiles
    const uint32_t segi = blockIdx.x * kGroups + grp, nseg = gridDim.x * kGroups;
    const uint32_t tile_first = uint32_t(uint64_t(segi) * ntiles / nseg);
    const uint32_t tile_last = uint32_t(uint64_t(segi + 1) * ntiles / nseg);
    const int64_t seg_begin = int64_t(a.pos_begin) + int64_t(tile_first) * kTile;
    const int64_t seg_end = min(int64_t(a.pos_end), int64_t(a.pos_begin) + int64_t(tile_last) * kTile);
    PhaseRec *seg = a.seg + size_t(segi) * a.seg_cap;
    uint32_t seg_n = 0;                            // uniform across the group
    unsigned long long my_preambles = 0;

    // The groups of a CTA start together; a.skew puts them out of step so that one group's dense phase A overlaps
    // the sparse phases B..D of the others: 1 = groups 1.. start a fraction of a tile time late (costs no issue
    // slots), 2 = a shorter first tile for every group but the first (one extra tile per group), 0 = nothing.
    int64_t tile_base = seg_begin;
#ifdef ADSBCU_PHASE_TIMING
    long long t_prev = 0;
#endif
    int first_len = a.skew == 2 ? kTile - ((kTile / kGroups * grp) & ~7) : kTile;
    if (a.skew == 1 && grp) __nanosleep(a.skew_ns * grp);
    if (a.skew >= 10 && grp >= int(a.skew - 10)) tile_base = seg_end;   // timing experiment only: idle groups
    // a thread's 24 samples + the 4 samples of lookahead its correlators need: 3 x 16 B + 8 B of I/Q
    uint4 raw[3];
    uint2 rawx;
    // tile-local sample indices [lo, hi) that exist in the feed: everything else comes from the carry / zero padding
    auto valid_range = [&](int64_t x0t, int &lo, int &hi) {
        lo = int(max(int64_t(0), -x0t));
        hi = int(min(int64_t(kTileSamples + 8), n - x0t));
    };
    auto load_raw = [&](int64_t base) {             // the tile whose first start position is `base`
        const int64_t x0t = base - kOverlap - kTileLead;
        int lo, hi;
        valid_range(x0t, lo, hi);
        const int i = kRun * tid;
        // magnitudes-in mode (SC16 inputs were converted to uint16 magnitudes): also 2 bytes per sample
        const uint8_t *p = (kMagIn ? reinterpret_cast<const uint8_t *>(a.mag_in) : a.iq) + 2 * (x0t + i);
#pragma unroll
        for (int r = 0; r < 3; ++r)
            if (i + 8 * r >= lo && i + 8 * r + 8 <= hi) raw[r] = __ldg(reinterpret_cast<const uint4 *>(p + 16 * r));
        if (i + 24 >= lo && i + 28 <= hi) rawx = __ldg(reinterpret_cast<const uint2 *>(p + 48));
    };
    // magnitude of feed sample xe on the slow path (stream edges): carry, scalar I/Q, or zero padding
    auto mag_edge = [&](int64_t xe) -> uint32_t {
        if (xe < 0) return (xe + kOverlap >= 0) ? uint32_t(a.carry_mag[xe + kOverlap]) : 0u;
        if (xe < n) return kMagIn ? uint32_t(a.mag_in[xe]) : mag_iq(s_sq, a.iq[2 * xe], a.iq[2 * xe + 1]);
        return 0u;
    };
    if (tile_base < seg_end) load_raw(tile_base);
    while (tile_base < seg_end) {
        group_sync(grp);   // previous tile fully consumed
        PHASE_MARK(1);
        const int npos = int(min(int64_t(first_len), seg_end - tile_base));   // start positions in this tile
        first_len = kTile;
        const int64_t next_base = tile_base + npos;
        // shared index i <-> feed sample x = x0 + i ; start position s (tile-local) sits at i = s + kTileLead
        const int64_t x0 = tile_base - kOverlap - kTileLead;

        // ---- phase A: magnitudes, rising/falling edge bits and all slicer decisions of 24 samples -------
        {
            const int64_t x = x0 + kRun * tid;
            const int i0 = kRun * tid;
            int v_lo, v_hi;
            valid_range(x0, v_lo, v_hi);
            uint32_t P[14];                          // P[k] = {mag[2k], mag[2k+1]} of this thread's run, k = 12, 13: lookahead
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const int64_t xv = x + 8 * r;
                if (i0 + 8 * r >= v_lo && i0 + 8 * r + 8 <= v_hi) {
                    const uint4 d = raw[r];
                    const uint32_t c = central_bits(d.x) & central_bits(d.y) & central_bits(d.z) & central_bits(d.w);
                    if (kMagIn) {
                        P[4 * r] = d.x; P[4 * r + 1] = d.y; P[4 * r + 2] = d.z; P[4 * r + 3] = d.w;
                    } else if ((c & 0x80808080u) == 0x80808080u) {
                        P[4 * r] = lut_pair(lut, d.x); P[4 * r + 1] = lut_pair(lut, d.y);
                        P[4 * r + 2] = lut_pair(lut, d.z); P[4 * r + 3] = lut_pair(lut, d.w);
                    } else {
                        P[4 * r] = mag_pair_any(s_sq, d.x); P[4 * r + 1] = mag_pair_any(s_sq, d.y);
                        P[4 * r + 2] = mag_pair_any(s_sq, d.z); P[4 * r + 3] = mag_pair_any(s_sq, d.w);
                    }
                } else {
#pragma unroll
                    for (int e = 0; e < 4; ++e) P[4 * r + e] = mag_edge(xv + 2 * e) | (mag_edge(xv + 2 * e + 1) << 16);
                }
                *reinterpret_cast<uint4 *>(&s_mag[kRun * tid + 8 * r]) = make_uint4(P[4 * r], P[4 * r + 1], P[4 * r + 2], P[4 * r + 3]);
            }
            if (i0 + 24 >= v_lo && i0 + 28 <= v_hi) {
                const uint32_t c = central_bits(rawx.x) & central_bits(rawx.y);
                if (kMagIn) { P[12] = rawx.x; P[13] = rawx.y; }
                else if ((c & 0x80808080u) == 0x80808080u) { P[12] = lut_pair(lut, rawx.x); P[13] = lut_pair(lut, rawx.y); }
                else { P[12] = mag_pair_any(s_sq, rawx.x); P[13] = mag_pair_any(s_sq, rawx.y); }
            } else {
                P[12] = mag_edge(x + 24) | (mag_edge(x + 25) << 16);
                P[13] = mag_edge(x + 26) | (mag_edge(x + 27) << 16);
            }
            // first differences nd[e] = mag[e+1] - mag[e] give everything else:
            //   edge bits      mag[e] < mag[e+1]  <=>  -nd[e] < 0 ;  mag[e] > mag[e+1]  <=>  nd[e] < 0
            //   correlators    slice_sub(&m[e]) = (5-sub) d0 + (2+sub) d1 (+ d2 for sub 4), d = -nd  (demod_2400.c:45-59);
            //                  their NEGATION is accumulated so that "correlation > 0" is the sign bit
            constexpr uint32_t kUp = w2(-1, 1);
            auto nd_at = [&](int e) -> int {         // e: compile-time after unrolling
                const uint32_t X = (e & 1) ? __funnelshift_r(P[e >> 1], P[(e >> 1) + 1], 16) : P[e >> 1];
                return dp2a_lo(X, kUp, 0);
            };
            uint32_t acc[12], L = 0, G = 0;
#pragma unroll
            for (int c = 0; c < 12; ++c) acc[c] = 0;
            // Two samples per round share the multiples of the middle difference: with a = nd[e], b = nd[e+1],
            // c = nd[e+2], d = nd[e+3] the negated correlators are
            //   sample e   : (5-sub) a + (2+sub) b (+ c for sub 4)  = a * (5-sub) + K[2+sub]
            //   sample e+1 : (5-sub) b + (2+sub) c (+ d for sub 4)  = c * (2+sub) + K[5-sub]        K[k] = k b
            // i.e. one multiply-add each on top of the five multiples K[2..6]
            int na = nd_at(0);
#pragma unroll
            for (int e = 0; e < kRun; e += 2) {
                const int nb = nd_at(e + 1), nc = nd_at(e + 2), ndd = nd_at(e + 3);
                G = __funnelshift_l(uint32_t(na), G, 1);
                L = __funnelshift_l(uint32_t(-na), L, 1);
                G = __funnelshift_l(uint32_t(nb), G, 1);
                L = __funnelshift_l(uint32_t(-nb), L, 1);
                const int K[7] = {0, nb, 2 * nb, 3 * nb, 4 * nb, 5 * nb, 6 * nb};
                int v[10];
#pragma unroll
                for (int sub = 0; sub < 5; ++sub) {
                    v[sub] = na * (5 - sub) + K[2 + sub];
                    v[5 + sub] = nc * (2 + sub) + K[5 - sub];
                }
                v[4] += nc;
                v[9] += ndd;
#pragma unroll
                for (int q = 0; q < 10; ++q) {
                    const int c = (5 * e + q) % 12;
                    acc[c] = __funnelshift_l(uint32_t(v[q]), acc[c], 1);
                }
                na = nc;
            }
#pragma unroll
            for (int c = 0; c < 12; ++c) gs.plane[c][tid] = uint16_t(acc[c]);
            L = __brev(L) >> 8; G = __brev(G) >> 8;   // bit e <-> sample e of the run
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                gs.u.edge.lbits[3 * tid + r] = uint8_t(L >> (8 * r));
                gs.u.edge.gbits[3 * tid + r] = uint8_t(G >> (8 * r));
            }
        }
        if (tid == 0) { gs.npre = 0; gs.ncand = 0; gs.nkeep = 0; gs.nq[0] = 0; gs.nq[1] = 0; }
        if (next_base < seg_end) load_raw(next_base);                   // in flight during phases B..D
        group_sync(grp);
        PHASE_MARK(2);

        // ---- phase B: preamble edge patterns, 32 start positions per thread ----------------------
        if (tid < kPosWords) {
            const uint32_t *Lw = reinterpret_cast<const uint32_t *>(gs.u.edge.lbits), *Gw = reinterpret_cast<const uint32_t *>(gs.u.edge.gbits);
            const uint32_t l0 = Lw[tid], g0 = Gw[tid];
            const uint32_t l1 = Lw[tid + 1], l2 = Lw[tid + 2], g1 = Gw[tid + 1], g2 = Gw[tid + 2];
            // bit b of myL/myG <-> start position 32*tid + b (smem index + kTileLead)
            const uint32_t myL = funnel_r(l0, l1, kTileLead), nxL = funnel_r(l1, l2, kTileLead);
            const uint32_t myG = funnel_r(g0, g1, kTileLead), nxG = funnel_r(g1, g2, kTileLead);
            const uint32_t L0 = myL, L2 = funnel_r(myL, nxL, 2), L3 = funnel_r(myL, nxL, 3), L8 = funnel_r(myL, nxL, 8),
                           L9 = funnel_r(myL, nxL, 9), L10 = funnel_r(myL, nxL, 10), L11 = funnel_r(myL, nxL, 11);
            const uint32_t G1 = funnel_r(myG, nxG, 1), G2 = funnel_r(myG, nxG, 2), G3 = funnel_r(myG, nxG, 3),
                           G4 = funnel_r(myG, nxG, 4), G9 = funnel_r(myG, nxG, 9), G10 = funnel_r(myG, nxG, 10),
                           G12 = funnel_r(myG, nxG, 12);
            const uint32_t early = G1 & L2, late = G10 & L11 & L9 & G4;
            const uint32_t pat = (early & G3 & L8 & G9 & (L10 | L11))          // phase 3, 4
                               | (early & G4 & L8 & G10 & L11)                 // phase 5
                               | (late & L3 & (G1 | G2));                      // phase 6, 7
            uint32_t cand = pat & L0 & G12;
            const int first = 32 * tid;
            if (first + 32 > npos) cand = first >= npos ? 0u : cand & (0xFFFFFFFFu >> (32 - (npos - first)));
            while (cand) {
                const int b = __ffs(cand) - 1;
                cand &= cand - 1;
                const uint32_t slot = atomicAdd(&gs.npre, 1u);
                if (slot < kPreCap) gs.u.edge.pre[slot] = uint16_t(first + b); else a.ctr->overflow = 1u;
            }
        }
        group_sync(grp);
        PHASE_MARK(3);
        {   // full preamble test (signal/noise ratio, quiet zones) on the compacted list
            const uint32_t npre = min(gs.npre, uint32_t(kPreCap));
            for (uint32_t i = tid; i < npre; i += kGroupThreads) {
                const uint32_t pos = gs.u.edge.pre[i];
                if (preamble_ok(&s_mag[pos + kTileLead])) {
                    const uint32_t slot = atomicAdd(&gs.ncand, 1u);
                    if (slot < kCandCap) gs.cand[slot] = uint16_t(pos); else a.ctr->overflow = 1u;
                    ++my_preambles;
                }
            }
        }
        group_sync(grp);
        PHASE_MARK(4);

        // ---- phase C: messages are bit strings of the symbol planes ----------------------------------
        // C1 reads the DF field (5 bits) of every (preamble, phase) and queues the items the DF pre-filter
        //    keeps (demod_2400.c:317-331);
        // C2 one item per thread: 56 / 112 plane bits, CRC-24 syndromes, repair, classification.
        const uint32_t ncand = min(gs.ncand, uint32_t(kCandCap));
        for (uint32_t c0 = 0; c0 < ncand; c0 += kCandChunk) {
            if (c0) {                                // (the first round's counters were cleared before phase B)
                if (tid < 2) gs.nq[tid] = 0;
                group_sync(grp);
            }
            const uint32_t nitems = min(uint32_t(kCandChunk), ncand - c0) * 5u;
            for (uint32_t it = tid; it < nitems; it += kGroupThreads) {
                const uint32_t pos = gs.cand[c0 + it / 5u];
                const uint32_t ph = 4u + it % 5u;
                const uint32_t t0 = 5u * (pos + kTileLead + 19u) + ph;
                const uint32_t c = t0 % 12u, n0 = t0 / 12u, f = n0 / 10u, r = n0 - 10u * f;
                const uint32_t w20 = (uint32_t(gs.plane[c][f]) << 10) | gs.plane[c][f + 1];
                const uint32_t dfbit = 1u << ((w20 >> (15u - r)) & 31u);
                const uint32_t item = pos | ((ph - 4u) << 16);
                if (valid_long & dfbit) gs.queue[atomicAdd(&gs.nq[0], 1u)] = item;
                else if (valid_short & dfbit) gs.queue[kQueueCap - 1 - atomicAdd(&gs.nq[1], 1u)] = item;
            }
            group_sync(grp);
            PHASE_MARK(5);
            const uint32_t nlong = gs.nq[0], nshort = gs.nq[1];
            for (uint32_t it = tid; it < nlong + nshort; it += kGroupThreads) {
                const bool is_long = it < nlong;
                const uint32_t item = is_long ? gs.queue[it] : gs.queue[kQueueCap - 1 - (it - nlong)];
                const uint32_t pos = item & 0xFFFFu, ph = 4u + (item >> 16);
                const uint32_t t0 = 5u * (pos + kTileLead + 19u) + ph;
                const uint32_t c = t0 % 12u, n0 = t0 / 12u;
                // the 10-bit fields from n0 on, repacked into a dense bit stream (every field is < 1024, so the
                // shifted fields do not overlap and add up), then shifted to the message's first bit
                const uint16_t *pf = gs.plane[c] + n0 / 10u;
                const uint32_t r = n0 % 10u;
                const uint32_t f3 = pf[3], f6 = pf[6], f9 = pf[9];
                const uint32_t w0 = (uint32_t(pf[0]) << 22) + (uint32_t(pf[1]) << 12) + (uint32_t(pf[2]) << 2) + (f3 >> 8);
                const uint32_t w1 = (f3 << 24) + (uint32_t(pf[4]) << 14) + (uint32_t(pf[5]) << 4) + (f6 >> 6);
                const uint32_t w2 = (f6 << 26) + (uint32_t(pf[7]) << 16) + (uint32_t(pf[8]) << 6) + (f9 >> 4);
                uint32_t M[4] = {__funnelshift_l(w1, w0, r), __funnelshift_l(w2, w1, r), 0u, 0u};
                uint32_t xs = 0, xl = 0;
                if (is_long) {
                    const uint32_t f12 = pf[12];
                    const uint32_t w3 = (f9 << 28) + (uint32_t(pf[10]) << 18) + (uint32_t(pf[11]) << 8) + (f12 >> 2);
                    M[2] = __funnelshift_l(w3, w2, r);
                    M[3] = __funnelshift_l(f12 << 30, w3, r) & 0xFFFF0000u;
                    xl = crc_long(M, s_crcb);
                    // the 56-bit syndrome is also needed when a short DF was sliced long (DF0/DF5 are within two
                    // bit flips of DF17/18, so --fix-2bit slices them to 14 bytes but scores them as 56-bit A/P)
                    if (!(M[0] >> 31) || (fixable_short & (1u << (M[0] >> 27)))) xs = crc_short(M, s_crcb);
                } else {
                    M[1] &= 0xFFFFFF00u;
                    xs = crc_short(M, s_crcb);
                }
                const Classified cl = classify(frame_from_words(M, is_long), xs, xl, a.tab, sm.occ_long, sm.occ_short);
                if (!cl.keep) continue;
                const uint32_t bi = bitmap_index(cl.test_key);
                if (cl.kind == K_DF17_0 || cl.kind == K_DF11_ACQ_0 || (cl.kind == K_DF18_0 && !df18_non_icao(cl.fixed)))
                    atomicOr(&a.bitmap[bi >> 5], 1u << (bi & 31));
                const uint32_t slot = atomicAdd(&gs.nkeep, 1u);
                if (slot < kKeepCap)
                    gs.keep[slot] = make_uint2((pos << 3) | (ph - 4u), bi | ((ph - 4u) << 25) | (cl.kind << 28));
                else
                    a.ctr->overflow = 1u;
            }
        }

        // ---- phase D: ordered emission of this tile's records into the segment ----------------
        // counting sort by 64-position bucket, then rank inside the (tiny) bucket
        group_sync(grp);
        PHASE_MARK(6);
        const uint32_t nkeep = min(gs.nkeep, uint32_t(kKeepCap));
        uint32_t *hist = gs.hist, *start = hist + kBuckets, *fill = hist + 2 * kBuckets;
        uint2 *sorted = gs.u.sorted;
        static_assert((kTileSamples << 3) <= (kBuckets << kBucketShift), "phase D buckets");
        if (seg_n + nkeep > a.seg_cap) {
            if (tid == 0) a.ctr->overflow = 1u;
        } else if (nkeep) {
            for (int i = tid; i < kBuckets; i += kGroupThreads) { hist[i] = 0; fill[i] = 0; }
            group_sync(grp);
            PHASE_MARK(7);
            for (uint32_t i = tid; i < nkeep; i += kGroupThreads) atomicAdd(&hist[gs.keep[i].x >> kBucketShift], 1u);
            group_sync(grp);
            PHASE_MARK(7);
            if (tid < 32) {
                constexpr int kPer = kBuckets / 32;
                uint32_t c[kPer], sum = 0;
#pragma unroll
                for (int k = 0; k < kPer; ++k) { c[k] = hist[kPer * tid + k]; sum += c[k]; }
                uint32_t incl = sum;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (tid >= d) incl += t; }
                uint32_t run = incl - sum;
#pragma unroll
                for (int k = 0; k < kPer; ++k) { start[kPer * tid + k] = run; run += c[k]; }
            }
            group_sync(grp);
            PHASE_MARK(7);
            for (uint32_t i = tid; i < nkeep; i += kGroupThreads) {
                const uint2 me = gs.keep[i];
                sorted[start[me.x >> kBucketShift] + atomicAdd(&fill[me.x >> kBucketShift], 1u)] = me;
            }
            group_sync(grp);
            PHASE_MARK(7);
            for (uint32_t i = tid; i < nkeep; i += kGroupThreads) {
                const uint2 me = sorted[i];
                const uint32_t lo = start[me.x >> kBucketShift], hi = lo + hist[me.x >> kBucketShift];
                uint32_t rank = lo;
                for (uint32_t j = lo; j < hi; ++j) rank += sorted[j].x < me.x;       // keys are unique
                PhaseRec r;
                r.pos = uint32_t(tile_base) + (me.x >> 3);
                r.info = me.y;
                seg[seg_n + rank] = r;
            }
            seg_n += nkeep;
        }
        tile_base = next_base;
    }
    if (tid == 0) {
        a.seg_count[segi] = seg_n;
        if (seg_n) atomicAdd(&a.ctr->n_cand, seg_n);
    }
    my_preambles = __reduce_add_sync(0xFFFFFFFFu, uint32_t(my_preambles));
    if (lane == 0 && my_preambles) atomicAdd(&a.ctr->preambles, my_preambles);
}

// Per segment, in place, order preserving.  A record is dropped when it can never be the accepted phase
// of its position:
//  * its score needs a known address (mode_s.c:370-446) and the address is not in the bitmap of addresses
//    that may ever enter the ICAO filter (mode_s.c:781-788);
//  * another phase of the same position with the SAME filter key beats it whether or not that key is known
//    (higher score_rank, or equal rank and lower phase: demod_2400.c:336 keeps the first best).
__device__ __forceinline__ bool kind_needs_known(uint32_t kind)
{
    return rank_of(kind, false) < kAcceptThreshold;
}
__global__ void __launch_bounds__(256) filter_kernel(ScanArgs a)
{
    __shared__ uint32_t s_warp[8];
    __shared__ PhaseRec s_r[4 + 256 + 4];
    PhaseRec *seg = a.seg + size_t(blockIdx.x) * a.seg_cap;
    const uint32_t n = min(a.seg_count[blockIdx.x], a.seg_cap);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const PhaseRec none{0xFFFFFFFFu, 0u};
    if (tid < 4) s_r[tid] = none;
    uint32_t w = 0;                                                  // uniform
    for (uint32_t base = 0; base < n; base += 256) {
        const uint32_t i = base + tid;
        s_r[4 + tid] = i < n ? seg[i] : none;
        if (tid < 4) s_r[260 + tid] = base + 256 + tid < n ? seg[base + 256 + tid] : none;
        __syncthreads();                                             // all reads of this batch are done
        const PhaseRec r = s_r[4 + tid];
        bool keep = false;
        if (i < n) {
            const uint32_t key = r.info & 0x1FFFFFFu, kind = r.info >> 28, ph = (r.info >> 25) & 7u;
            const bool unknown = kind_needs_known(kind) && !((a.bitmap[key >> 5] >> (key & 31)) & 1u);
            keep = true;
            const int r0 = rank_of(kind, false), r1 = rank_of(kind, true);
#pragma unroll
            for (int d = -4; d <= 4; ++d) {
                if (d == 0) continue;
                const PhaseRec q = s_r[4 + tid + d];
                if (q.pos != r.pos || (q.info & 0x1FFFFFFu) != key) continue;
                const uint32_t qk = q.info >> 28, qp = (q.info >> 25) & 7u;
                const int q0 = rank_of(qk, false), q1 = rank_of(qk, true);
                if ((q0 > r0 || (q0 == r0 && qp < ph)) && (q1 > r1 || (q1 == r1 && qp < ph))) keep = false;
            }
            if (keep && unknown) {
                keep = false;
                if (a.deferred) {
                    const uint32_t slot = atomicAdd(a.n_deferred, 1u);
                    if (slot < a.deferred_cap) a.deferred[slot] = r; else a.ctr->overflow = 1u;
                }
            }
        }
        const uint32_t bal = __ballot_sync(0xFFFFFFFFu, keep);
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        uint32_t before = 0, total = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) { const uint32_t c = s_warp[k]; if (k < warp) before += c; total += c; }
        if (keep) seg[w + before + __popc(bal & ((1u << lane) - 1u))] = r;   // index <= i: never ahead of the reads
        w += total;
        if (tid < 4) s_r[tid] = s_r[256 + tid];                      // look-back for the next batch
        __syncthreads();
    }
    if (tid == 0) a.seg_count[blockIdx.x] = w;
}

// Sharded captures: after the address sets of the other shards were OR-ed into the bitmap, the deferred
// records whose address is in it now are laid out flat over the segments for evaluate_kernel.
__global__ void __launch_bounds__(256) revive_kernel(ScanArgs a)
{
    const uint32_t n = min(*a.n_deferred, a.deferred_cap);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const PhaseRec r = a.deferred[i];
        const uint32_t key = r.info & 0x1FFFFFFu;
        if (!((a.bitmap[key >> 5] >> (key & 31)) & 1u)) continue;
        const uint32_t slot = atomicAdd(&a.ctr->n_cand, 1u);
        if (slot < a.nseg * a.seg_cap) a.seg[slot] = r; else a.ctr->overflow = 1u;
    }
}
__global__ void revive_counts_kernel(ScanArgs a)
{
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= a.nseg) return;
    const uint64_t total = min(uint64_t(a.ctr->n_cand), uint64_t(a.nseg) * a.seg_cap), lo = uint64_t(g) * a.seg_cap;
    a.seg_count[g] = total > lo ? uint32_t(min(uint64_t(a.seg_cap), total - lo)) : 0u;
}
__global__ void or_words_kernel(uint32_t *dst, const uint32_t *src, uint32_t n)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) dst[i] |= src[i];
}

// One warp per surviving record, one CTA per segment.  The record's 290 samples are converted
// to magnitudes into a per-warp shared buffer, lane l slices message bits l, l+32, l+64, l+96,
// ballots assemble the frame, the CRC-24 syndromes are an XOR reduction of per-bit syndromes.
__global__ void __launch_bounds__(256) evaluate_kernel(ScanArgs a)
{
    __shared__ uint16_t s_m[8][320];
    __shared__ __align__(16) FullRec s_rec[8];
    __shared__ uint32_t s_part[8];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // exclusive prefix of the survivor counts of the segments before this one
    uint32_t part = 0;
    for (uint32_t g = tid; g < blockIdx.x; g += 256) part += a.seg_count[g];
    part = __reduce_add_sync(0xFFFFFFFFu, part);
    if (lane == 0) s_part[warp] = part;
    __syncthreads();
    uint32_t offset = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) offset += s_part[k];
    const uint32_t nrec = min(a.seg_count[blockIdx.x], a.seg_cap);
    if (blockIdx.x == gridDim.x - 1 && tid == 0) {
        a.ctr->n_full = offset + nrec;
        if (offset + nrec > a.full_cap) a.ctr->overflow = 1u;
    }
    const PhaseRec *seg = a.seg + size_t(blockIdx.x) * a.seg_cap;
    const int64_t n = a.nsamples;
    const float *sq = a.tab->sq;
    const uint32_t *bit_synd = a.tab->bit_synd;
    uint16_t *m = s_m[warp];

    for (uint32_t r = warp; r < nrec; r += 8) {
        const uint32_t o = offset + r;
        if (o >= a.full_cap) break;
        const PhaseRec pr = seg[r];
        const int ph = 4 + int((pr.info >> 25) & 7u);
        const int64_t xb = int64_t(pr.pos) - kOverlap + 19;           // feed sample of m[j+19]
        __syncwarp();
        for (int i = lane; i < 320; i += 32) {
            const int64_t x = xb + i;
            uint32_t v = 0;
            if (x < 0) v = (x + kOverlap >= 0) ? a.carry_mag[x + kOverlap] : 0u;
            else if (x < n) v = a.mag_in ? uint32_t(a.mag_in[x]) : mag_iq(sq, a.iq[2 * x], a.iq[2 * x + 1]);
            m[i] = uint16_t(v);
        }
        __syncwarp();
        uint32_t M[4], xs = 0, xl = 0;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            const int k = 32 * w + lane;
            const int u = ph + 12 * k, s = u / 5, sub = u - 5 * s;
            const bool bit = k < 112 && slice_corr(m[s], m[s + 1], m[s + 2], m[s + 3], sub) > 0;
            M[w] = __brev(__ballot_sync(0xFFFFFFFFu, bit));
        }
        const uint32_t dfbit = 1u << (M[0] >> 27);
        const bool is_long = (a.tab->valid_long & dfbit) != 0;
        if (!is_long && !(a.tab->valid_short & dfbit)) continue;      // cannot happen for a record the scan kept
        if (is_long) { M[3] &= 0xFFFF0000u; } else { M[1] &= 0xFFFFFF00u; M[2] = 0u; M[3] = 0u; }
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            const int k = 32 * w + lane;
            if ((M[w] >> (31 - lane)) & 1u) {
                xl ^= bit_synd[k < 112 ? k : 0];
                if (k < 56) xs ^= bit_synd[k + 56];
            }
        }
        xl = __reduce_xor_sync(0xFFFFFFFFu, xl);
        xs = __reduce_xor_sync(0xFFFFFFFFu, xs);
        const Frame f = frame_from_words(M, is_long);
        const Classified c = classify(f, xs, xl, a.tab);              // uniform across the warp
        // signal power window (demod_2400.c:382-389): length from the *uncorrected* DF
        const int siglen = (frame_df(f) & 16u) ? 268 : 134;
        unsigned long long acc = 0;
        for (int i = lane; i < siglen; i += 32) { const uint32_t v = m[i]; acc += uint64_t(v * v); }
#pragma unroll
        for (int d = 16; d; d >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, d);
        if (lane == 0) {
            FullRec &q = s_rec[warp];
            q.pos = pr.pos;
            q.phase = uint8_t(ph);
            q.kind = uint8_t(c.keep ? c.kind : K_COUNT);
            q.corrections = int8_t(c.corrections < 0 ? 0 : c.corrections);
            q.msgtype = uint8_t(frame_df(c.fixed));
#pragma unroll
            for (int b = 0; b < 14; ++b) {
                q.verbatim[b] = uint8_t(frame_byte(f, b));
                q.msg[b] = uint8_t(frame_byte(c.fixed, b));
            }
            q.crc = c.crc;
            q.test_key = c.test_key;
            q.add_key = kUnchecked;
            const uint32_t aa = frame_aa(c.fixed);
            switch (q.msgtype) {
            case 11: q.addr = aa; if (c.kind == K_DF11_ACQ_0) q.add_key = aa; break;
            case 17: q.addr = aa; if (c.kind == K_DF17_0) q.add_key = aa; break;
            case 18:
                q.addr = aa | (df18_non_icao(c.fixed) ? kNonIcao : 0u);
                if (c.kind == K_DF18_0) q.add_key = q.addr | kNtFlag;
                break;
            default: q.addr = c.crc; break;
            }
            // mm.signalLevel (demod_2400.c:382-403): IEEE double divisions, the same on the device as on the host
            q.signal_level = double(acc) / 65535.0 / 65535.0 / double(siglen);
        }
        __syncwarp();
        if (lane < 4) reinterpret_cast<uint4 *>(&a.full[o])[lane] = reinterpret_cast<const uint4 *>(&s_rec[warp])[lane];
    }
}

// mag_lut[lut_offset(I, Q) / 2] = magnitude(I, Q), computed exactly as magnitude_kernel does
__global__ void build_lut_kernel(DevTables *tab)
{
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;    // I | Q << 8
    const uint32_t i = idx & 255u, q = idx >> 8;
    if (idx < 65536u && lut_covers(i) && lut_covers(q)) tab->mag_lut[lut_offset(i, q) >> 1] = uint16_t(mag_iq(tab->sq, i, q));
}

// SC16 / SC16Q11 input converters (convert.c:46-92 -> dsp/impl/magnitude_sc16.c:27-45, magnitude_sc16q11.c:29-47,
// the "exact_float" implementations the generic flavour registers first): |I|,|Q| scaled by 2 (or 32), float
// squares and sum each rounded to nearest (no FMA contraction), sqrtf, clamp, truncate.
__global__ void __launch_bounds__(256) convert_sc16_kernel(const int16_t *in, uint32_t nsamples, uint16_t *mag, int scale)
{
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < nsamples; k += gridDim.x * blockDim.x) {
        const int2 raw = make_int2(in[2 * k], in[2 * k + 1]);
        const float i = float(abs(raw.x) * scale), q = float(abs(raw.y) * scale);
        float m = __fsqrt_rn(__fadd_rn(__fmul_rn(i, i), __fmul_rn(q, q)));
        if (m > 65535.0f) m = 65535.0f;
        mag[k] = uint16_t(__float2uint_rz(m));
    }
}

__global__ void carry_mag_kernel(const uint16_t *mag_in, uint32_t nsamples, uint16_t *carry_mag)
{
    const int i = threadIdx.x;
    uint32_t v = 0;
    if (i < kOverlap) {
        const int64_t x = int64_t(nsamples) - kOverlap + i;
        v = x >= 0 ? mag_in[x] : carry_mag[x + kOverlap];
    }
    __syncthreads();
    if (i < kOverlap) carry_mag[i] = uint16_t(v);
}

__global__ void carry_kernel(const uint8_t *iq, uint32_t nsamples, uint16_t *carry_mag, const DevTables *tab)
{
    // new_carry[i] = magnitude of stream sample (end - 326 + i); samples before this feed come from the old carry
    const int i = threadIdx.x;
    uint32_t v = 0;
    if (i < kOverlap) {
        const int64_t x = int64_t(nsamples) - kOverlap + i;
        if (x >= 0) v = mag_iq(tab->sq, iq[2 * x], iq[2 * x + 1]);
        else v = carry_mag[x + kOverlap];
    }
    __syncthreads();
    if (i < kOverlap) carry_mag[i] = uint16_t(v);
}

__global__ void __launch_bounds__(256) magnitude_kernel(const uint8_t *iq, uint32_t nsamples, uint16_t *mag, const DevTables *tab)
{
    __shared__ float s_sq[256];
    s_sq[threadIdx.x] = tab->sq[threadIdx.x];
    __syncthreads();
    const uint32_t nvec = nsamples / 8;
    for (uint32_t v = blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += gridDim.x * blockDim.x) {
        const uint4 d = __ldg(reinterpret_cast<const uint4 *>(iq) + v);
        const uint32_t w[4] = {d.x, d.y, d.z, d.w};
        uint32_t o[4];
#pragma unroll
        for (int e = 0; e < 4; ++e)
            o[e] = mag_iq(s_sq, w[e] & 0xFF, (w[e] >> 8) & 0xFF) | (mag_iq(s_sq, (w[e] >> 16) & 0xFF, w[e] >> 24) << 16);
        reinterpret_cast<uint4 *>(mag)[v] = make_uint4(o[0], o[1], o[2], o[3]);
    }
    if (blockIdx.x == 0)
        for (uint32_t k = nvec * 8 + threadIdx.x; k < nsamples; k += blockDim.x)
            mag[k] = uint16_t(mag_iq(s_sq, iq[2 * k], iq[2 * k + 1]));
}

}  // namespace

#ifdef ADSBCU_PHASE_TIMING
extern "C" void adsbcu_debug_phase_cycles(unsigned long long *out, int reset)
{
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_phase_cycles, sizeof(g_phase_cycles));
    if (reset) { unsigned long long z[8] = {}; cudaMemcpyToSymbol(g_phase_cycles, z, sizeof(z)); }
}
#endif

cudaError_t init_kernels()
{
    const cudaError_t e = cudaFuncSetAttribute(scan_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(ScanSmem)));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(scan_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(ScanSmem)));
}

cudaError_t launch_scan(const ScanArgs &a, cudaStream_t s)
{
    if (a.pos_end <= a.pos_begin || !a.nseg) return cudaSuccess;
    if (a.mag_in) ADSBCU_LAUNCH(scan_kernel<true>, a.nseg / kGroups, kGroups * kGroupThreads, sizeof(ScanSmem), s, a);
    else ADSBCU_LAUNCH(scan_kernel<false>, a.nseg / kGroups, kGroups * kGroupThreads, sizeof(ScanSmem), s, a);
    return cudaGetLastError();
}

cudaError_t launch_build_lut(DevTables *tab, cudaStream_t s)
{
    ADSBCU_LAUNCH(build_lut_kernel, 256, 256, 0, s, tab);
    return cudaGetLastError();
}

cudaError_t launch_filter(const ScanArgs &a, cudaStream_t s)
{
    if (a.pos_end <= a.pos_begin || !a.nseg) return cudaSuccess;
    ADSBCU_LAUNCH(filter_kernel, a.nseg, 256, 0, s, a);
    return cudaGetLastError();
}

cudaError_t launch_evaluate(const ScanArgs &a, cudaStream_t s)
{
    if (a.pos_end <= a.pos_begin || !a.nseg) return cudaSuccess;
    ADSBCU_LAUNCH(evaluate_kernel, a.nseg, 256, 0, s, a);
    return cudaGetLastError();
}

cudaError_t launch_revive(const ScanArgs &a, cudaStream_t s)
{
    ADSBCU_LAUNCH(revive_kernel, 296, 256, 0, s, a);
    ADSBCU_LAUNCH(revive_counts_kernel, (a.nseg + 255) / 256, 256, 0, s, a);
    ADSBCU_LAUNCH(evaluate_kernel, a.nseg, 256, 0, s, a);
    return cudaGetLastError();
}

// Chunk completion without a host call: the counters go to mapped host memory, the sequence number last (the
// evaluate kernel's records were written to mapped memory earlier in the same stream).
__global__ void signal_kernel(const ChunkCounters *d, volatile ChunkCounters *h, uint32_t seq)
{
    h->preambles = d->preambles; h->n_cand = d->n_cand; h->n_full = d->n_full; h->overflow = d->overflow;
    __threadfence_system();
    h->pad = seq;
}
cudaError_t launch_signal(const ChunkCounters *d_ctr, ChunkCounters *h_ctr_dev, uint32_t seq, cudaStream_t s)
{
    ADSBCU_LAUNCH(signal_kernel, 1, 1, 0, s, d_ctr, h_ctr_dev, seq);
    return cudaGetLastError();
}

cudaError_t launch_or_words(uint32_t *dst, const uint32_t *src, uint32_t nwords, cudaStream_t s)
{
    ADSBCU_LAUNCH(or_words_kernel, 296, 256, 0, s, dst, src, nwords);
    return cudaGetLastError();
}

cudaError_t launch_convert_sc16(const int16_t *in, uint32_t nsamples, uint16_t *mag, int q11, cudaStream_t s)
{
    if (!nsamples) return cudaSuccess;
    ADSBCU_LAUNCH(convert_sc16_kernel, 148 * 8, 256, 0, s, in, nsamples, mag, q11 ? 32 : 2);
    return cudaGetLastError();
}

cudaError_t launch_carry_mag(const uint16_t *mag_in, uint32_t nsamples, uint16_t *carry_mag, cudaStream_t s)
{
    ADSBCU_LAUNCH(carry_mag_kernel, 1, 352, 0, s, mag_in, nsamples, carry_mag);
    return cudaGetLastError();
}

cudaError_t launch_carry(const uint8_t *iq, uint32_t nsamples, uint16_t *carry_mag, const DevTables *tab, cudaStream_t s)
{
    ADSBCU_LAUNCH(carry_kernel, 1, 352, 0, s, iq, nsamples, carry_mag, tab);
    return cudaGetLastError();
}

cudaError_t launch_magnitude(const uint8_t *iq, uint32_t nsamples, uint16_t *mag, const DevTables *tab, cudaStream_t s)
{
    if (!nsamples) return cudaSuccess;
    ADSBCU_LAUNCH(magnitude_kernel, 148 * 8, 256, 0, s, iq, nsamples, mag, tab);
    return cudaGetLastError();
}

}  // namespace adsbcu
