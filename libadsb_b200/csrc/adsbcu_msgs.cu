// message_kernel -- the per-(preamble, phase) half of the 1090 MHz path (sm_100a).
//
// Input: the start positions the front kernel confirmed and its symbol planes.  For every position and each of the
// five try_phases (demod_2400.c:223-347) the message is 56 / 112 consecutive bits of one plane, so this kernel never
// touches a magnitude: it reads the DF field, applies the DF pre-filter (demod_2400.c:317-331), repacks the bits,
// forms both CRC-24 syndromes from per-byte tables in shared memory (crc.c:65-81), runs correctMessage and the
// state-independent half of scoreModesMessage (mode_s.c:252-466) and keeps what can still be accepted:
//   * records that are acceptable whatever the ICAO filter holds, and Address/Parity-style records whose address
//     is already in the "may ever be in the filter" set, are written out as evaluated records (FullRec);
//   * records whose address has not been announced yet are parked as 8-byte (position, info) entries; the order
//     kernel looks at them again when the whole chunk's announcements are known.
// The work list is flat: a "round" is up to 256 preambles of one front-kernel segment, the CTAs of a fixed grid take
// rounds in turn, and inside a round every warp handles its 32 preambles on its own (ballot-compacted work queues,
// no CTA barrier), so that many independent warps hide the latency of the plane reads.
#include "adsbcu_device.cuh"

namespace adsbcu {
namespace {

#ifndef ADSBCU_MSG_MINCTAS            // resident CTAs per SM the register allocation aims for (measurement knob)
#define ADSBCU_MSG_MINCTAS 6
#endif
constexpr int kMThreads = 256;
constexpr int kMWarps = kMThreads / 32;
constexpr int kMRound = kMThreads;             // preambles per round: 32 per warp
constexpr int kMaxSeg = 160;                   // front-kernel CTAs (one per SM)

struct MsgSmem {
    uint32_t crcb[11 * 256];                   // DevTables::crc_byte
    uint32_t occ_long[kSyndSlotsLong / 32], occ_short[kSyndSlotsShort / 32];
    uint32_t queue[kMWarps][160];              // per warp: 112-bit items from the front, 56-bit items from the back: lane | (phase-4) << 8
    uint32_t slow[kMWarps][160];               // per warp: items that need the general classification
    PhaseRec park[kMWarps][160];               // per warp: the round's parked records (one list reservation per round)
    uint32_t round_first[kMaxSeg + 1];         // first round of every segment
};

__global__ void __launch_bounds__(kMThreads, ADSBCU_MSG_MINCTAS) message_kernel(ChunkArgs a)
{
    ADSBCU_DYN_SMEM(smem_raw);
    MsgSmem &sm = *reinterpret_cast<MsgSmem *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    {
        const uint4 *src = reinterpret_cast<const uint4 *>(&a.tab->crc_byte[0][0]);
        uint4 *dst = reinterpret_cast<uint4 *>(sm.crcb);
        for (int i = tid; i < 11 * 256 / 4; i += kMThreads) dst[i] = src[i];
    }
    for (int i = tid; i < kSyndSlotsLong / 32; i += kMThreads) sm.occ_long[i] = a.tab->occ_long[i];
    for (int i = tid; i < kSyndSlotsShort / 32; i += kMThreads) sm.occ_short[i] = a.tab->occ_short[i];
    // rounds of every segment: exclusive prefix of ceil(count / 256) over the (at most kMaxSeg) segments
    if (tid < kMaxSeg + 1) sm.round_first[tid] = tid < int(a.nseg) ? (min(a.pre_count[tid], a.pre_cap) + kMRound - 1) / kMRound : 0u;
    const uint32_t valid_long = a.tab->valid_long, valid_short = a.tab->valid_short, fixable_short = a.tab->fixable_short,
                   fixable_long = a.tab->fixable_long;
    __syncthreads();
    if (tid < 32) {                            // one warp: each lane scans kMaxSeg / 32 consecutive counts
        constexpr int kPer = (kMaxSeg + 32) / 32;
        uint32_t c[kPer], sum = 0;
#pragma unroll
        for (int k = 0; k < kPer; ++k) { const int i = kPer * tid + k; c[k] = i <= kMaxSeg ? sm.round_first[i] : 0u; sum += c[k]; }
        uint32_t incl = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (tid >= d) incl += t; }
        uint32_t run = incl - sum;
#pragma unroll
        for (int k = 0; k < kPer; ++k) { const int i = kPer * tid + k; if (i <= kMaxSeg) sm.round_first[i] = run; run += c[k]; }
    }
    __syncthreads();
    const uint32_t nrounds = sm.round_first[a.nseg];
    uint32_t *queue = sm.queue[warp];

    // From here on every warp works alone: 32 preambles of the round, no CTA barrier.
    for (uint32_t round = blockIdx.x; round < nrounds; round += gridDim.x) {
        uint32_t seg = 0;                      // last segment whose first round is <= round and that has rounds
        for (uint32_t step = 128; step; step >>= 1)
            if (seg + step < a.nseg && sm.round_first[seg + step] <= round) seg += step;
        const uint32_t c0 = (round - sm.round_first[seg]) * kMRound + 32u * warp;
        const uint32_t npre = min(a.pre_count[seg], a.pre_cap);
        if (c0 >= npre) continue;              // warp-uniform
        const bool have = c0 + lane < npre;
        const uint32_t mypos = have ? a.pre[size_t(seg) * a.pre_cap + c0 + lane] : 0u;     // chunk-local position
        // C1: the DF field (first five bits) of the lane's preamble at each try_phase decides 112 bits, 56 bits or nothing
        uint32_t nlong = 0, nshort = 0;
#pragma unroll
        for (uint32_t k = 0; k < 5u; ++k) {
            const uint32_t t0 = 5u * (mypos + kFLead + 19u) + 4u + k;
            const uint32_t c = t0 % 12u, n0 = t0 / 12u, f = n0 / 10u, r = n0 - 10u * f;
            const uint16_t *pf = a.planes + size_t(c) * a.plane_stride + f;
            const uint32_t w20 = have ? (uint32_t(pf[0]) << 10) | pf[1] : 0u;
            const uint32_t dfbit = 1u << ((w20 >> (15u - r)) & 31u);
            const bool is_l = have && (valid_long & dfbit), is_s = have && !is_l && (valid_short & dfbit);
            const uint32_t bl = __ballot_sync(0xFFFFFFFFu, is_l), bs = __ballot_sync(0xFFFFFFFFu, is_s);
            const uint32_t below = (1u << lane) - 1u;
            if (is_l) queue[nlong + __popc(bl & below)] = uint32_t(lane) | (k << 8);
            if (is_s) queue[159u - nshort - __popc(bs & below)] = uint32_t(lane) | (k << 8);
            nlong += __popc(bl);
            nshort += __popc(bs);
        }
        __syncwarp();
        // parked records collect in the warp's buffer; the round ends with one reservation in the segment's list
        uint32_t npark = 0;
        auto park_records = [&](bool park, const PhaseRec &parked) {
            const uint32_t bp = __ballot_sync(0xFFFFFFFFu, park);
            if (park) sm.park[warp][npark + __popc(bp & ((1u << lane) - 1u))] = parked;
            npark += __popc(bp);
        };
        // C2: one item per lane and pass
        uint32_t nslow = 0;
        for (uint32_t it0 = 0; it0 < nlong + nshort; it0 += 32) {
            const uint32_t it = it0 + lane;
            const bool live = it < nlong + nshort;
            const bool is_long = it < nlong;
            const uint32_t item = live ? (is_long ? queue[it] : queue[159u - (it - nlong)]) : 0u;
            const uint32_t cpos = __shfl_sync(0xFFFFFFFFu, mypos, int(item & 31u)), ph = 4u + (item >> 8);
            bool park = false, slow = false;
            PhaseRec parked{0u, 0u};
            if (live) {
                const uint32_t t0 = 5u * (cpos + kFLead + 19u) + ph;
                const uint32_t c = t0 % 12u, n0 = t0 / 12u;
                uint32_t M[4], xs, xl, key = 0;
                slice_planes(a.planes + size_t(c) * a.plane_stride + n0 / 10u, n0 % 10u, is_long, sm.crcb, fixable_short, M, xs, xl);
                const uint32_t pos = a.pos_begin + cpos;                              // feed-relative position
                // the common case first (noise: no repair applies), the general classification only where it does not
                int verdict = classify_fast(M[0], M[0] == 0u && (M[1] >> 8) == 0u, xs, xl, fixable_long, fixable_short, sm.occ_long, sm.occ_short, key);
                if (verdict == 1) {
                    // acceptable only for a known address: parked; the order kernel evaluates the ones whose address has
                    // been announced by the end of the chunk (no look at the address set here: one memory round trip less)
                    park = true;
                    parked = PhaseRec{pos, bitmap_index(key) | ((ph - 4u) << 25) | (uint32_t(K_AP) << 28)};
                }
                slow = verdict == 2;
            }
            // the rare general case is queued: it runs after the passes, several lanes at a time instead of one
            const uint32_t bq = __ballot_sync(0xFFFFFFFFu, slow);
            if (slow) sm.slow[warp][nslow + __popc(bq & ((1u << lane) - 1u))] = item | (is_long ? 0x10000u : 0u);
            nslow += __popc(bq);
            park_records(park, parked);
        }
        __syncwarp();
        // the queued items: slicing again is cheaper than carrying 112 bits per lane through the passes
        for (uint32_t s0 = 0; s0 < nslow; s0 += 32) {
            const bool live = s0 + lane < nslow;
            const uint32_t item = live ? sm.slow[warp][s0 + lane] : 0u;
            const bool is_long = (item & 0x10000u) != 0u;
            const uint32_t cpos = __shfl_sync(0xFFFFFFFFu, mypos, int(item & 31u)), ph = 4u + ((item >> 8) & 7u);
            bool park = false;
            PhaseRec parked{0u, 0u};
            if (live) {
                const uint32_t t0 = 5u * (cpos + kFLead + 19u) + ph;
                const uint32_t c = t0 % 12u, n0 = t0 / 12u;
                uint32_t M[4], xs, xl;
                slice_planes(a.planes + size_t(c) * a.plane_stride + n0 / 10u, n0 % 10u, is_long, sm.crcb, fixable_short, M, xs, xl);
                const uint32_t pos = a.pos_begin + cpos;
                const Frame f = frame_from_words(M, is_long);
                const Classified cl = classify(f, xs, xl, a.tab, sm.occ_long, sm.occ_short);
                if (cl.keep) {
                    const uint32_t bi = bitmap_index(cl.test_key);
                    if (cl.kind == K_DF17_0 || cl.kind == K_DF11_ACQ_0 || (cl.kind == K_DF18_0 && !df18_non_icao(cl.fixed)))
                        atomicOr(&a.bitmap[bi >> 5], 1u << (bi & 31));
                    if (rank_of(cl.kind, false) < kAcceptThreshold && !((a.bitmap[bi >> 5] >> (bi & 31)) & 1u)) {
                        park = true;
                        parked = PhaseRec{pos, bi | ((ph - 4u) << 25) | (cl.kind << 28)};
                    } else {
                        const uint32_t slot = atomicAdd(&a.rec_count[seg], 1u);
                        if (slot < a.rec_cap) fill_rec(a.rec[size_t(seg) * a.rec_cap + slot], pos, ph, f, cl);
                        else a.ctr->overflow = 1u;
                    }
                }
            }
            park_records(park, parked);
        }
        __syncwarp();
        if (npark) {
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(&a.pend_count[seg], npark);
            base = __shfl_sync(0xFFFFFFFFu, base, 0);
            if (base + npark > a.pend_cap) { if (lane == 0) a.ctr->overflow = 1u; }
            else for (uint32_t i = lane; i < npark; i += 32) a.pend[size_t(seg) * a.pend_cap + base + i] = sm.park[warp][i];
        }
        __syncwarp();
    }
}

}  // namespace

cudaError_t init_messages()
{
    return cudaFuncSetAttribute(message_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(MsgSmem)));
}

// grid: a fixed number of CTAs that take rounds in turn; rec_count / pend_count must be zero on entry
cudaError_t launch_messages(const ChunkArgs &a, uint32_t grid, cudaStream_t s)
{
    if (a.pos_end <= a.pos_begin || !a.nseg) return cudaSuccess;
    if (a.nseg > uint32_t(kMaxSeg)) return cudaErrorInvalidValue;
    ADSBCU_LAUNCH(message_kernel, grid, kMThreads, sizeof(MsgSmem), s, a);
    return cudaGetLastError();
}

}  // namespace adsbcu
