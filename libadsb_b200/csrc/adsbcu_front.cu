// front_kernel -- the dense half of the 1090 MHz path (sm_100a).
//
//   uint8 I/Q  ->  magnitudes (128 KiB LUT gather from shared memory, bank-swizzled so that near-DC samples spread over
//                  the banks)
//              ->  every slicer decision S(t) of the stream, 5 per sample, written once as 12 "symbol planes"
//              ->  falling-edge bits -> bit-parallel preamble edge patterns -> the reference's full preamble test
//              ->  list of start positions that pass (demod_2400.c:155-218)
//
// Replaces starch_magnitude_uc8 (magnitude_uc8.c:30-63), the slicers (demod_2400.c:45-59) and the preamble detector
// (demod_2400.c:134-218).  Everything that is per (preamble, phase) -- DF pre-filter, CRC, repair, scoring -- is the
// message kernel (adsbcu_msgs.cu), which only READS bits of the planes written here.
//
// Symbol planes: with "time" t = 5 * sample + sub (sub = correlator phase 0..4) the decision S(t) = [slice_sub(&m[sample]) > 0]
// depends on t only, and symbol k of a message whose first data symbol is at t0 sits at t0 + 12 k.  A message is
// therefore 56 / 112 consecutive bits of plane (t0 mod 12).  A lane owns 24 consecutive samples = 120 values of t =
// 10 bits of each of the 12 planes; it writes them as twelve 10-bit fields planes[c][run].
//
// Geometry: one CTA of 32 warps per SM (minus the SMs the host keeps free for the order/resolve kernels); the warps
// share the LUT and nothing else.  A warp works through tiles of 768 samples (= 768 start positions; the 24
// magnitudes of look-ahead the preamble test needs are three extra vector gathers) entirely on its own -- magnitudes
// in a warp-private shared buffer, edge bits exchanged by shuffle, candidates compacted with a warp scan -- so the
// steady state has no CTA-wide barrier at all and the dense phase of one warp overlaps the sparse phases of others.
#include "adsbcu_device.cuh"

namespace adsbcu {
namespace {

constexpr int kPreW = 256;                      // candidate positions a warp lists per tile (more: per-lane loop, no overflow)

struct FrontWarp {
    __align__(16) uint16_t mag[kWTile + kFHalo + 8];
    uint16_t pre[kPreW];
};
struct FrontSmem {
    __align__(16) uint16_t lut[65536];          // DevTables::mag_lut_full
    FrontWarp w[kFWarps];
    uint32_t nconf;                             // confirmed preambles of this CTA's segment so far
};
static_assert(sizeof(FrontSmem) <= 227 * 1024, "front kernel shared memory exceeds the sm_100 per-CTA limit");

// magnitudes of the two samples {I0,Q0,I1,Q1} of a 32-bit word: the LUT index keeps I and Q where they are in the
// 16-bit sample and xors three low bits of Q into index bits 3..5 (lutf_index), one shift + one LOP3 for both samples
__device__ __forceinline__ void lut_two(const uint16_t *lut, uint32_t w, int &m0, int &m1)
{
    const uint32_t u = w ^ ((w >> 5) & 0x00380038u);
    m0 = lut[u & 0xFFFFu];
    m1 = lut[u >> 16];
}

// The reference's preamble test at p[0..18] (demod_2400.c:155-218), branch-free: the phase-3..7 patterns are an
// if / else-if chain there, so the first pattern that matches decides high / base_signal / base_noise.
__device__ __forceinline__ bool preamble_full(const uint16_t *p)
{
    const int p0 = p[0], p1 = p[1], p2 = p[2], p3 = p[3], p4 = p[4], p5 = p[5], p6 = p[6], p7 = p[7], p8 = p[8], p9 = p[9],
              p10 = p[10], p11 = p[11], p12 = p[12], p13 = p[13], p14 = p[14], p15 = p[15], p16 = p[16], p17 = p[17], p18 = p[18];
    const bool g12 = p1 > p2, l23 = p2 < p3, g23 = p2 > p3, g34 = p3 > p4, l34 = p3 < p4, g45 = p4 > p5, l89 = p8 < p9,
               g910 = p9 > p10, l910 = p9 < p10, l1011 = p10 < p11, g1011 = p10 > p11, l1112 = p11 < p12;
    const bool a = g12 & l23 & g34 & l89 & g910;
    const bool pat3 = a & l1011;
    const bool pat4 = a & l1112;
    const bool pat5 = g12 & l23 & g45 & l89 & g1011 & l1112;
    const bool b = l34 & g45 & l910 & g1011 & l1112;
    const bool pat6 = g12 & b;
    const bool pat7 = g23 & b;
    // sums of the first matching pattern
    int hsum, sig, noise;
    if (pat3)      { hsum = p1 + p3 + p9 + p11 + p12;      sig = p1 + p3 + p9;        noise = p5 + p6 + p7; }
    else if (pat4) { hsum = p1 + p3 + p9 + p12;            sig = p1 + p3 + p9 + p12;  noise = p5 + p6 + p7 + p8; }
    else if (pat5) { hsum = p1 + p3 + p4 + p9 + p10 + p12; sig = p1 + p12;            noise = p6 + p7; }
    else if (pat6) { hsum = p1 + p4 + p10 + p12;           sig = p1 + p4 + p10 + p12; noise = p5 + p6 + p7 + p8; }
    else           { hsum = p1 + p2 + p4 + p10 + p12;      sig = p4 + p10 + p12;      noise = p6 + p7 + p8; }
    const int high = hsum >> 2;                 // (sum) / 4 of non-negative ints
    const int quiet = max(max(max(p5, p6), max(p7, p8)), max(max(p14, p15), max(max(p16, p17), p18)));
    return (p0 < p1) & (p12 > p13) & (pat3 | pat4 | pat5 | pat6 | pat7) & (2 * sig >= 3 * noise) & (quiet < high);
}

// Slicer decisions and falling-edge bits of the eight samples E0 .. E0+7 of a lane's run; m[k] is the magnitude of
// sample E0 + k (k < 11).  First differences nd[e] = mag[e+1] - mag[e] give everything:
//   falling edge   mag[e] > mag[e+1]  <=>  nd[e] < 0
//   correlators    slice_sub(&m[e]) = -((5-sub) nd[e] + (2+sub) nd[e+1] (+ nd[e+2] for sub 4))   (demod_2400.c:45-59);
//                  the NEGATED value is formed so that "correlation > 0" is its sign bit; the five of one sample are
//                  v0 = 5 nd[e] + 2 nd[e+1], then + step, + step, + step, + step + nd[e+2] with step = nd[e+1] - nd[e]
template <int E0>
__device__ __forceinline__ void decide8(const int (&m)[12], uint32_t (&acc)[12], uint32_t &G)
{
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int na = m[k + 1] - m[k], nb = m[k + 2] - m[k + 1], nc = m[k + 3] - m[k + 2];
        G = __funnelshift_l(uint32_t(na), G, 1);
        const int step = nb - na;
        int v[5];
        v[0] = 5 * na + 2 * nb;
        v[1] = v[0] + step;
        v[2] = v[1] + step;
        v[3] = v[2] + step;
        v[4] = v[3] + step + nc;
#pragma unroll
        for (int q = 0; q < 5; ++q) {
            constexpr int dummy = 0;
            (void)dummy;
            const int c = (5 * (E0 + k) + q) % 12;
            acc[c] = __funnelshift_l(uint32_t(v[q]), acc[c], 1);   // (acc << 1) | (correlation > 0)
        }
    }
}

// kMagIn: the input is uint16 magnitudes (ChunkArgs::mag_in, SC16 front end) instead of uint8 I/Q
template <bool kMagIn>
__global__ void __launch_bounds__(kFWarps * 32, 1) front_kernel(ChunkArgs a)
{
    ADSBCU_DYN_SMEM(smem_raw);
    FrontSmem &sm = *reinterpret_cast<FrontSmem *>(smem_raw);

    // ---- CTA prologue: the 128 KiB magnitude LUT into shared memory (TMA bulk copies on one mbarrier) --------
    if (threadIdx.x == 0) sm.nconf = 0;
    if (!kMagIn) {
#ifdef ADSBCU_EMU
        for (uint32_t i = threadIdx.x; i < 65536u; i += blockDim.x) sm.lut[i] = a.tab->mag_lut_full[i];
#else
        __shared__ __align__(8) unsigned long long s_bar;
        const uint32_t bar = uint32_t(__cvta_generic_to_shared(&s_bar));
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            constexpr uint32_t kLutBytes = 65536u * 2u;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(kLutBytes) : "memory");
            for (uint32_t o = 0; o < kLutBytes; o += 32768u)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(uint32_t(__cvta_generic_to_shared(sm.lut)) + o),
                               "l"(reinterpret_cast<const unsigned char *>(a.tab->mag_lut_full) + o), "r"(32768u), "r"(bar) : "memory");
        }
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(bar) : "memory");
#endif
    }
    __syncthreads();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    FrontWarp &ws = sm.w[warp];
    uint16_t *s_mag = ws.mag;
    const uint16_t *lut = sm.lut;

    // this CTA's segment: a contiguous run of the chunk's warp-tiles, dealt round-robin to its warps
    const uint32_t npos_chunk = a.pos_end - a.pos_begin;
    const uint32_t ntiles = (npos_chunk + kFNeed + kWTile - 1) / kWTile;
    const uint32_t segi = blockIdx.x, nseg = gridDim.x;
    const uint32_t tile_first = uint32_t(uint64_t(segi) * ntiles / nseg);
    const uint32_t tile_last = uint32_t(uint64_t(segi + 1) * ntiles / nseg);
    uint32_t *pre_out = a.pre + size_t(segi) * a.pre_cap;
    const int64_t nvis = a.nsamples;               // feed samples this chunk may read

    uint4 raw[3];
    uint32_t rawh = 0;
    int nx_lo = 0, nx_hi = 0;                      // staged indices [lo, hi) of the prefetched tile that exist in the feed
    // staged index j of a tile <-> feed sample x0 + j
    const int64_t x0_base = int64_t(a.pos_begin) - kOverlap - kFLead;
    auto x0_of = [&](uint32_t tile) { return x0_base + int64_t(tile) * kWTile; };
    const uint8_t *src = kMagIn ? reinterpret_cast<const uint8_t *>(a.mag_in) : a.iq;   // both 2 bytes per sample
    const int j0 = kFRun * lane;
    auto load_raw = [&](uint32_t tile) {
        const int64_t x0 = x0_of(tile);
        nx_lo = int(max(int64_t(0), -x0));
        nx_hi = int(min(int64_t(kWTile + kFHalo + 8), nvis - x0));
        const uint8_t *p = src + 2 * (x0 + j0);
#pragma unroll
        for (int r = 0; r < 3; ++r)
            if (j0 + 8 * r >= nx_lo && j0 + 8 * r + 8 <= nx_hi) raw[r] = __ldcs(reinterpret_cast<const uint4 *>(p + 16 * r));
        // the 24 samples of look-ahead past the tile: one pair per lane, lanes 0..11
        const int jh = kWTile + 2 * lane;
        if (lane < 12 && jh >= nx_lo && jh + 2 <= nx_hi) rawh = __ldcs(reinterpret_cast<const uint32_t *>(src + 2 * (x0 + jh)));
    };
    // magnitude of feed sample x on the slow path (stream edges): carry, one sample, or zero padding
    auto mag_edge = [&](int64_t x) -> int {
        if (x < 0) return (x + kOverlap >= 0) ? int(a.carry_mag[x + kOverlap]) : 0;
        if (x < nvis) return kMagIn ? int(a.mag_in[x]) : int(lut[lutf_index(uint32_t(a.iq[2 * x]) | (uint32_t(a.iq[2 * x + 1]) << 8))]);
        return 0;
    };
    auto two = [&](uint32_t w, int &m0, int &m1) {
        if (kMagIn) { m0 = int(w & 0xFFFFu); m1 = int(w >> 16); } else lut_two(lut, w, m0, m1);
    };
    auto pack = [](int lo, int hi) { return uint32_t(lo) | (uint32_t(hi) << 16); };

    uint32_t tile = tile_first + warp;
    if (tile < tile_last) load_raw(tile);
    for (; tile < tile_last; tile += kFWarps) {
        const int64_t x0 = x0_of(tile);
        const uint32_t tile_pos = tile * kWTile;           // chunk-local position of staged index kFLead
        const int npos = int(min(int64_t(kWTile), int64_t(npos_chunk) - int64_t(tile_pos)));   // <= 0: decisions only
        const int v_lo = nx_lo, v_hi = nx_hi;
        const bool interior = v_lo == 0 && v_hi == kWTile + kFHalo + 8;       // whole tile inside the feed (warp-uniform)

        // ---- phase A: magnitudes, falling-edge bits and all slicer decisions of 24 samples per lane ---------
        uint32_t acc[12], G = 0;
#pragma unroll
        for (int c = 0; c < 12; ++c) acc[c] = 0;
        {
            int m[12];
            // vector r of the lane: samples 8r .. 8r+7
            auto vec8 = [&](int r, const uint4 &d, int (&o)[8]) {
                if (interior || (j0 + 8 * r >= v_lo && j0 + 8 * r + 8 <= v_hi)) {
                    two(d.x, o[0], o[1]); two(d.y, o[2], o[3]); two(d.z, o[4], o[5]); two(d.w, o[6], o[7]);
                } else {
#pragma unroll
                    for (int e = 0; e < 8; ++e) o[e] = mag_edge(x0 + j0 + 8 * r + e);
                }
                *reinterpret_cast<uint4 *>(&s_mag[j0 + 8 * r]) = make_uint4(pack(o[0], o[1]), pack(o[2], o[3]), pack(o[4], o[5]), pack(o[6], o[7]));
            };
            int v0[8], v1[8], v2[8];
            vec8(0, raw[0], v0);
            vec8(1, raw[1], v1);
            // look-ahead past the tile (lanes 0..11, one pair each): the preamble test of the tile's last start positions
            // needs it, and so do the last three correlators of lane 31
            int h0 = 0, h1 = 0;
            if (lane < 12) {
                const int jh = kWTile + 2 * lane;
                if (interior || (jh >= v_lo && jh + 2 <= v_hi)) two(rawh, h0, h1);
                else { h0 = mag_edge(x0 + jh); h1 = mag_edge(x0 + jh + 1); }
                *reinterpret_cast<uint32_t *>(&s_mag[jh]) = pack(h0, h1);
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) m[k] = v0[k];
#pragma unroll
            for (int k = 0; k < 4; ++k) m[8 + k] = v1[k];
            decide8<0>(m, acc, G);
            vec8(2, raw[2], v2);
#pragma unroll
            for (int k = 0; k < 8; ++k) m[k] = v1[k];
#pragma unroll
            for (int k = 0; k < 4; ++k) m[8 + k] = v2[k];
            decide8<8>(m, acc, G);
            // the three magnitudes after the lane's run: the next lane's first three, or the look-ahead for lane 31
            int n0 = __shfl_down_sync(0xFFFFFFFFu, v0[0], 1), n1 = __shfl_down_sync(0xFFFFFFFFu, v0[1], 1), n2 = __shfl_down_sync(0xFFFFFFFFu, v0[2], 1);
            const int e0 = __shfl_sync(0xFFFFFFFFu, h0, 0), e1 = __shfl_sync(0xFFFFFFFFu, h1, 0), e2 = __shfl_sync(0xFFFFFFFFu, h0, 1);
            if (lane == 31) { n0 = e0; n1 = e1; n2 = e2; }
#pragma unroll
            for (int k = 0; k < 8; ++k) m[k] = v2[k];
            m[8] = n0; m[9] = n1; m[10] = n2; m[11] = 0;
            decide8<16>(m, acc, G);
        }
        {
            const size_t run = size_t(tile) * 32 + lane;
#pragma unroll
            for (int c = 0; c < 12; ++c) a.planes[size_t(c) * a.plane_stride + run] = uint16_t(acc[c]);
        }
        G = __brev(G) >> 8;                                 // bit e <-> sample e of the lane's run
        if (tile + kFWarps < tile_last) load_raw(tile + kFWarps);   // in flight during the sparse phases
        __syncwarp();                                        // the tile's magnitudes are in shared memory

        if (npos > 0) {
            // ---- phase B: preamble edge patterns, 24 start positions per lane ---------------------------------
            // Only "greater" bits are kept; "less" is taken as "not greater", a superset on ties that the exact test
            // below removes again.
            uint32_t nextG = __shfl_down_sync(0xFFFFFFFFu, G, 1);
            // the bits past the tile come from the staged look-ahead: one comparison per lane, one ballot
            const uint32_t haloG = __ballot_sync(0xFFFFFFFFu, lane < 16 && s_mag[kWTile + lane] > s_mag[kWTile + lane + 1]);
            if (lane == 31) nextG = haloG;
            // bit b of (lo, hi) <-> staged sample 24*lane + b; start position s of the lane sits at staged index s + kFLead
            const uint32_t lo = G | (nextG << 24), hi = nextG >> 8;
            const uint32_t myG = __funnelshift_r(lo, hi, kFLead), nxG = hi >> kFLead;
            const uint32_t myL = ~myG, nxL = ~nxG;
            auto Gs = [&](int s) { return __funnelshift_r(myG, nxG, s); };
            auto Ls = [&](int s) { return __funnelshift_r(myL, nxL, s); };
            const uint32_t L0 = myL, L2 = Ls(2), L3 = Ls(3), L8 = Ls(8), L9 = Ls(9), L10 = Ls(10), L11 = Ls(11);
            const uint32_t G1 = Gs(1), G2 = Gs(2), G3 = Gs(3), G4 = Gs(4), G9 = Gs(9), G10 = Gs(10), G12 = Gs(12);
            const uint32_t early = G1 & L2, late = G10 & L11 & L9 & G4;
            const uint32_t pat = (early & G3 & L8 & G9 & (L10 | L11))          // phase 3, 4
                               | (early & G4 & L8 & G10 & L11)                 // phase 5
                               | (late & L3 & (G1 | G2));                      // phase 6, 7
            uint32_t cand = pat & L0 & G12 & 0xFFFFFFu;
            const int first = kFRun * lane;
            if (first + kFRun > npos) cand = first >= npos ? 0u : cand & (0xFFFFFFu >> (kFRun - (npos - first)));
            // warp scan of the candidate counts -> compact list in shared memory
            const uint32_t cnt = __popc(cand);
            uint32_t incl = cnt;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
            const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 31);
            // ---- phase B2: the reference's full test (edges exactly, signal/noise ratio, quiet zones) ----------
            auto confirm = [&](bool ok, uint32_t pos) {     // warp-aggregated append to the segment's list
                const uint32_t bal = __ballot_sync(0xFFFFFFFFu, ok);
                if (!bal) return;
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(&sm.nconf, uint32_t(__popc(bal)));
                base = __shfl_sync(0xFFFFFFFFu, base, 0);
                if (ok) {
                    const uint32_t slot = base + __popc(bal & ((1u << lane) - 1u));
                    if (slot < a.pre_cap) pre_out[slot] = tile_pos + pos; else a.ctr->overflow = 1u;
                }
            };
            if (total <= uint32_t(kPreW)) {
                uint32_t slot = incl - cnt, cm = cand;
                while (cm) {
                    const int b = __ffs(cm) - 1;
                    cm &= cm - 1;
                    ws.pre[slot++] = uint16_t(first + b);
                }
                __syncwarp();
                for (uint32_t i0 = 0; i0 < total; i0 += 32) {
                    const uint32_t i = i0 + lane;
                    const uint32_t pos = i < total ? ws.pre[i] : 0u;
                    confirm(i < total && preamble_full(&s_mag[pos + kFLead]), pos);
                }
            } else {                                        // pathological density: every lane walks its own candidates
                uint32_t cm = cand;
                while (__any_sync(0xFFFFFFFFu, cm != 0u)) {
                    const int b = cm ? __ffs(cm) - 1 : 0;
                    const bool have = cm != 0u;
                    cm &= cm - 1;
                    confirm(have && preamble_full(&s_mag[first + b + kFLead]), uint32_t(first + b));
                }
            }
        }
        __syncwarp();                                        // everyone is done with mag[] and pre[] of this tile
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        a.pre_count[segi] = min(sm.nconf, a.pre_cap);
        if (sm.nconf) atomicAdd(&a.ctr->preambles, (unsigned long long)sm.nconf);
    }
}

// mag_lut_full[lutf_index(I | Q << 8)] = magnitude(I, Q) of the reference's table (tables.c:9-41), computed in registers
__global__ void build_lutf_kernel(DevTables *tab)
{
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;    // I | Q << 8
    if (s < 65536u) tab->mag_lut_full[lutf_index(s)] = uint16_t(mag_iq(tab->sq, s & 255u, s >> 8));
}

}  // namespace

cudaError_t init_front()
{
    const cudaError_t e = cudaFuncSetAttribute(front_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(FrontSmem)));
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(front_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(FrontSmem)));
}

cudaError_t launch_front(const ChunkArgs &a, cudaStream_t s)
{
    if (a.pos_end <= a.pos_begin || !a.nseg) return cudaSuccess;
    if (a.mag_in) ADSBCU_LAUNCH(front_kernel<true>, a.nseg, kFWarps * 32, sizeof(FrontSmem), s, a);
    else ADSBCU_LAUNCH(front_kernel<false>, a.nseg, kFWarps * 32, sizeof(FrontSmem), s, a);
    return cudaGetLastError();
}

cudaError_t launch_build_lutf(DevTables *tab, cudaStream_t s)
{
    ADSBCU_LAUNCH(build_lutf_kernel, 256, 256, 0, s, tab);
    return cudaGetLastError();
}

}  // namespace adsbcu
