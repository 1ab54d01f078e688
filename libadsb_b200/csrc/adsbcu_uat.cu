// Path C: 978 MHz UAT behind UAT978Handler::HandleData (reference UAT978.cpp:43-60).
//
// The reference maps every I/Q byte pair through a 64 Ki phase table (UAT978.cpp:52,76-100), stages the
// phases in a 65536-entry buffer and calls the dump978 C function process_buffer() (UAT978.cpp:9-10,54),
// which scans two bit alignments for the ADS-B / uplink sync words, slices frames around a per-frame
// centre frequency, Reed-Solomon-corrects them and hands them to dump_raw_message()
// (uat2json-wrapper.cpp:14).  dump978 itself is an absent submodule (github.com/ankurvdev/dump978 @
// a461b30): PARITY UNPINNED, the algorithm is restated from the published dump978 legacy demodulator
// and libfec (see oracle/oracle978.c for the same statement on the checker side).
//
//   uat_phase_kernel   I/Q -> phase (table gather), appended after the carried tail of the stream
//   uat_sync_kernel    every sample position: check_sync_word() for both sync words (mean dphi of the
//                      one / zero bits -> slicing centre -> at most 4 sync errors); candidates compacted
//   uat_decode_kernel  one thread per candidate: slice the frame, RS(48,34) / RS(30,18) or de-interleave
//                      + 6 x RS(92,72) (Berlekamp-Massey, Chien, Forney over GF(256)/0x187)
//   host pass          the order-dependent rest of process_buffer(): which alignment the scan tries,
//                      fewer-corrections-wins between index and index+1, skipping over a decoded frame,
//                      sync shift registers that keep stale bits across a skip, 65536-sample staging with
//                      carry-over exactly as UAT978.cpp:48-59 does it.
#include "../../include/adsbcu.h"
#include "adsbcu_internal.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

namespace adsbcu {

namespace {

constexpr int kSyncBits = 36;
constexpr unsigned long long kAdsbSync = 0xEACDDA4E2ull, kUplinkSync = 0x153225B1Dull;
constexpr int kLongBytes = 48, kShortBytes = 30, kLongData = 34, kShortData = 18;
constexpr int kUpBlocks = 6, kUpBlockBytes = 92, kUpBlockData = 72, kUpBytes = 552, kUpData = 432;
constexpr int kUpBits = kUpBytes * 8;
constexpr int kNeed = 2 * (kSyncBits + kUpBits) + 2;          // samples a candidate position may look at
constexpr int kTail = 9216;                                   // >= kNeed + sigma window lead, multiple of 8
constexpr int kSigmaLead = 8;                                 // sigma window of a record starts 8 samples before it

struct CandC { uint32_t x; int16_t center; uint16_t uplink; };   // x: index into the device phase buffer

struct RecC {                 // 480 bytes
    uint64_t s;               // absolute stream sample of the first sync bit
    uint32_t sigma[3];        // bit i: phi_difference(phi[s-8+i], phi[s-7+i]) > 0
    int16_t  rs_errors;
    uint8_t  uplink;          // 0 ADS-B, 1 ground uplink
    uint8_t  type;            // ADS-B: 1 basic, 2 long
    uint8_t  data[kUpData];
    uint8_t  pad[24];
};
static_assert(sizeof(RecC) == 480, "RecC layout");

struct CountersC { unsigned int n_cand, n_rec, overflow, pad; };

__device__ __forceinline__ int phi_diff(uint32_t from, uint32_t to)
{
    return int(int16_t(uint16_t(to - from)));                 // wrap to [-32768, 32767]
}

__global__ void __launch_bounds__(256) uat_phase_kernel(const uint8_t *iq, uint32_t n, const uint16_t *lut, uint16_t *phi)
{
    const uint16_t *iq16 = reinterpret_cast<const uint16_t *>(iq);   // {I,Q} little endian == table index (UAT978.cpp:88-96)
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) phi[i] = __ldg(&lut[iq16[i]]);
}

// positions [first, last) of the phase buffer; every position reads phi[x .. x+72)
__global__ void __launch_bounds__(256) uat_sync_kernel(const uint16_t *phi, uint32_t first, uint32_t last, CandC *cand, uint32_t cand_cap,
                                                       CountersC *ctr)
{
    __shared__ int16_t s_d[256 + 72];
    const uint32_t base = first + blockIdx.x * 256u;
    for (int i = threadIdx.x; i < 256 + 72; i += 256) {
        const uint32_t x = base + i;
        s_d[i] = x < last + 72 ? int16_t(phi_diff(phi[x], phi[x + 1])) : int16_t(0);
    }
    __syncthreads();
    const uint32_t x = base + threadIdx.x;
    if (x >= last) return;
    const int16_t *d = &s_d[threadIdx.x];
    int one_a = 0, zero_a = 0, one_u = 0, zero_u = 0;
#pragma unroll
    for (int i = 0; i < kSyncBits; ++i) {
        const int v = d[2 * i];
        if ((kAdsbSync >> (35 - i)) & 1ull) one_a += v; else zero_a += v;
        if ((kUplinkSync >> (35 - i)) & 1ull) one_u += v; else zero_u += v;
    }
    constexpr int kOnesA = __builtin_popcountll(kAdsbSync), kOnesU = __builtin_popcountll(kUplinkSync);
    const int ca = int(int16_t((one_a / kOnesA + zero_a / (kSyncBits - kOnesA)) / 2));
    const int cu = int(int16_t((one_u / kOnesU + zero_u / (kSyncBits - kOnesU)) / 2));
    int err_a = 0, err_u = 0;
#pragma unroll
    for (int i = 0; i < kSyncBits; ++i) {
        const int v = d[2 * i];
        err_a += ((kAdsbSync >> (35 - i)) & 1ull) ? (v < ca) : (v >= ca);
        err_u += ((kUplinkSync >> (35 - i)) & 1ull) ? (v < cu) : (v >= cu);
    }
    if (err_a <= 4) {
        const uint32_t slot = atomicAdd(&ctr->n_cand, 1u);
        if (slot < cand_cap) cand[slot] = CandC{x, int16_t(ca), 0}; else ctr->overflow = 1u;
    }
    if (err_u <= 4) {
        const uint32_t slot = atomicAdd(&ctr->n_cand, 1u);
        if (slot < cand_cap) cand[slot] = CandC{x, int16_t(cu), 1}; else ctr->overflow = 1u;
    }
}

// ---- GF(256) / Reed-Solomon (field polynomial 0x187, first root alpha^120, libfec char-codec behaviour) ----
struct Gf { uint8_t exp[256], log[256]; };                    // log[0] = 255 ("A0"), exp[255] = 0

__device__ __forceinline__ int mod255(int x) { while (x >= 255) { x -= 255; x = (x >> 8) + (x & 255); } return x; }

// In-place decode of a shortened code word of n = 255 - pad symbols with `nroots` parity symbols.
// Returns the number of error locations found, or -1 (block untouched) if the locator has fewer roots than
// its degree.  A location inside the virtual padding is counted but not applied, as libfec does.
// syndrome i of the block (libfec's Horner form): the O(n * nroots) part of the decoder, one syndrome per lane
__device__ __forceinline__ uint8_t rs_syndrome(const Gf &gf, const uint8_t *blk, int n, int i)
{
    constexpr int kFcr = 120;
    uint32_t s = blk[0];
    for (int j = 1; j < n; ++j) s = s == 0 ? blk[j] : blk[j] ^ gf.exp[mod255(gf.log[s] + kFcr + i)];
    return uint8_t(s);
}

// The rest of the decoder, in three warp-wide steps around a small per-block work area in shared memory: Berlekamp-Massey
// (one coefficient per lane), the Chien search -- 255 evaluations of the locator, the long serial part of libfec's decode_rs,
// each lane evaluating the polynomial directly at its own field elements -- and Forney (one root per lane).
struct RsWork {
    uint8_t syn[20];          // syndromes, log form
    uint8_t lam[21];          // error locator, log form
    uint8_t root[20], loc[20];
    uint8_t omg[21];          // error evaluator, log form
    int deg, count, nroots;
};

// Chien search by the whole warp: libfec walks i = 1..255 and stops at the deg-th root; here lane l evaluates
// lambda(alpha^i) for i = 32 r + l + 1 and the roots are collected in ascending i, the first deg of them.
__device__ void rs_chien_warp(const Gf &gf, RsWork &w, int lane)
{
    constexpr int kA0 = 255;
    const int deg = w.deg;
    int count = 0;
    for (int r = 0; r < 8 && count < deg; ++r) {
        const int i = 32 * r + lane + 1;
        bool is_root = false;
        if (i <= 255) {
            int q = 1;
            for (int j = deg; j > 0; --j)
                if (w.lam[j] != kA0) q ^= gf.exp[(int(w.lam[j]) + i * j) % 255];
            is_root = q == 0;
        }
        const uint32_t b = __ballot_sync(0xFFFFFFFFu, is_root);
        const int at = count + __popc(b & ((1u << lane) - 1u));
        if (is_root && at < deg) { w.root[at] = uint8_t(i); w.loc[at] = uint8_t(i - 1); }
        count += __popc(b);
    }
    __syncwarp();
    if (lane == 0) w.count = min(count, deg);
    __syncwarp();
}

// Berlekamp-Massey by the whole warp: lane j holds coefficient j of the locator (polynomial form) and of the auxiliary
// polynomial (log form); a round is one XOR reduction for the discrepancy and one shuffle for the shift.  Same arithmetic and
// same order of updates as libfec's decode_rs loop (restated in oracle/oracle978.c).  Returns 0 when all syndromes are zero.
__device__ int rs_bm_warp(const Gf &gf, RsWork &w, const uint8_t *syn_raw, int nroots, int lane)
{
    constexpr int kA0 = 255;
    const uint32_t raw = lane < nroots ? syn_raw[lane] : 0u;
    if (lane < nroots) w.syn[lane] = gf.log[raw];
    if (lane == 0) { w.nroots = nroots; w.deg = 0; w.count = 0; }
    const uint32_t any = __ballot_sync(0xFFFFFFFFu, raw != 0u);
    __syncwarp();
    if (!any) return 0;
    uint32_t lam = lane == 0 ? 1u : 0u, bb = lane == 0 ? 0u : uint32_t(kA0);
    int el = 0;
    for (int r = 1; r <= nroots; ++r) {
        uint32_t term = 0;
        if (lane < r && lam != 0u) {
            const uint32_t sy = w.syn[r - lane - 1];
            if (sy != uint32_t(kA0)) term = gf.exp[mod255(int(gf.log[lam]) + int(sy))];
        }
        const uint32_t disc = gf.log[__reduce_xor_sync(0xFFFFFFFFu, term)];
        uint32_t bprev = __shfl_up_sync(0xFFFFFFFFu, bb, 1);
        if (lane == 0) bprev = uint32_t(kA0);
        if (disc == uint32_t(kA0)) {
            bb = bprev;
        } else {
            const uint32_t tt = (lane >= 1 && bprev != uint32_t(kA0)) ? lam ^ gf.exp[mod255(int(disc) + int(bprev))] : lam;
            if (2 * el <= r - 1) {
                el = r - el;
                bb = lam == 0u ? uint32_t(kA0) : uint32_t(mod255(int(gf.log[lam]) - int(disc) + 255));
            } else {
                bb = bprev;
            }
            lam = tt;
        }
    }
    const uint32_t ll = lane <= nroots ? uint32_t(gf.log[lam]) : uint32_t(kA0);
    if (lane <= nroots) w.lam[lane] = uint8_t(ll);
    const uint32_t nz = __ballot_sync(0xFFFFFFFFu, ll != uint32_t(kA0));
    if (lane == 0) w.deg = nz ? 31 - __clz(int(nz)) : 0;
    __syncwarp();
    return 1;
}

// error evaluator and Forney by the whole warp: lane i forms omega_i, then lane j corrects the location of root j
__device__ int rs_forney_warp(const Gf &gf, RsWork &w, uint8_t *blk, int pad, int lane)
{
    constexpr int kA0 = 255, kFcr = 120;
    const int deg = w.deg, count = w.count, nroots = w.nroots;
    if (deg != count) return -1;
    const int dego = deg - 1;
    if (lane <= dego) {
        int t = 0;
        for (int j = lane; j >= 0; --j)
            if (w.syn[lane - j] != kA0 && w.lam[j] != kA0) t ^= gf.exp[mod255(w.syn[lane - j] + w.lam[j])];
        w.omg[lane] = gf.log[t];
    }
    __syncwarp();
    if (lane < count) {
        const int rt = w.root[lane];
        int num1 = 0, den = 0;
        for (int i = dego; i >= 0; --i)
            if (w.omg[i] != kA0) num1 ^= gf.exp[mod255(w.omg[i] + i * rt)];
        const int num2 = gf.exp[mod255(rt * (kFcr - 1) + 255)];
        for (int i = min(deg, nroots - 1) & ~1; i >= 0; i -= 2)
            if (w.lam[i + 1] != kA0) den ^= gf.exp[mod255(w.lam[i + 1] + i * rt)];
        if (num1 != 0 && w.loc[lane] >= pad) blk[w.loc[lane] - pad] ^= gf.exp[mod255(gf.log[num1] + gf.log[num2] + 255 - gf.log[den])];
    }
    __syncwarp();
    return count;
}

// one block, whole warp: -> corrections (or -1)
__device__ int rs_decode_warp(const Gf &gf, RsWork &w, uint8_t *blk, const uint8_t *syn_raw, int nroots, int pad, int lane)
{
    if (!rs_bm_warp(gf, w, syn_raw, nroots, lane)) return 0;
    rs_chien_warp(gf, w, lane);
    return rs_forney_warp(gf, w, blk, pad, lane);
}

__device__ __forceinline__ uint8_t slice_byte(const uint16_t *p, int center)
{
    uint32_t b = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) b = (b << 1) | uint32_t(phi_diff(p[2 * k], p[2 * k + 1]) > center);
    return uint8_t(b);
}

// One WARP per candidate (the per-call latency of the staging loop is the latency of its slowest candidate): the lanes
// slice the frame bytes and form the Reed-Solomon syndromes in parallel, the short algebraic part runs on one lane per
// RS block (six lanes for an uplink frame), the sign bits of the record come from ballots.
constexpr int kDecWarps = 4;
struct DecSmem {
    Gf gf;
    uint8_t buf[kDecWarps][kUpBlocks * kUpBlockBytes + 8];     // ADS-B: 48 frame bytes; uplink: six de-interleaved blocks of 92
    uint8_t syn[kDecWarps][kUpBlocks][20];
    int     res[kDecWarps][kUpBlocks];
    RsWork  work[kDecWarps][kUpBlocks];
};

__global__ void __launch_bounds__(kDecWarps * 32) uat_decode_kernel(const uint16_t *phi, uint64_t abs_of_index0, const CandC *cand, RecC *rec,
                                                                    uint32_t rec_cap, CountersC *ctr, const Gf *gf_g)
{
    __shared__ DecSmem sm;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) { sm.gf.exp[i] = gf_g->exp[i]; sm.gf.log[i] = gf_g->log[i]; }
    __syncthreads();
    const Gf &gf = sm.gf;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t *buf = sm.buf[warp];
    const uint32_t ncand = ctr->n_cand;
    for (uint32_t c = blockIdx.x * kDecWarps + warp; c < ncand; c += gridDim.x * kDecWarps) {
        const CandC cd = cand[c];
        const uint16_t *p = phi + cd.x + 2 * kSyncBits;
        int rs = 0, type = 0;
        bool ok = false;
        __syncwarp();
        if (!cd.uplink) {
            for (int b = lane; b < kLongBytes; b += 32) buf[b] = slice_byte(p + 16 * b, cd.center);
            __syncwarp();
            if (lane < 14) sm.syn[warp][0][lane] = rs_syndrome(gf, buf, kLongBytes, lane);          // long: RS(48,34)
            __syncwarp();
            int nfix = rs_decode_warp(gf, sm.work[warp][0], buf, sm.syn[warp][0], 14, 207, lane);
            if (nfix >= 0 && nfix <= 7 && (buf[0] >> 3) != 0) { type = 2; rs = nfix; ok = true; }
            else {                                                                                  // basic: RS(30,18) on the same bytes
                if (lane < 12) sm.syn[warp][0][lane] = rs_syndrome(gf, buf, kShortBytes, lane);
                __syncwarp();
                nfix = rs_decode_warp(gf, sm.work[warp][0], buf, sm.syn[warp][0], 12, 225, lane);
                if (nfix >= 0 && nfix <= 6 && (buf[0] >> 3) == 0) { type = 1; rs = nfix; ok = true; }
            }
        } else {
            // de-interleave: byte i of block b is frame byte 6 i + b
            for (int q = lane; q < kUpBytes; q += 32) buf[(q % kUpBlocks) * kUpBlockBytes + q / kUpBlocks] = slice_byte(p + 16 * q, cd.center);
            __syncwarp();
            for (int t = lane; t < kUpBlocks * 20; t += 32) sm.syn[warp][t / 20][t % 20] = rs_syndrome(gf, buf + (t / 20) * kUpBlockBytes, kUpBlockBytes, t % 20);
            __syncwarp();
            for (int blk = 0; blk < kUpBlocks; ++blk) {                                             // RS(92,72), block after block, warp-wide
                const int nfix = rs_decode_warp(gf, sm.work[warp][blk], buf + blk * kUpBlockBytes, sm.syn[warp][blk], 20, 163, lane);
                if (lane == 0) sm.res[warp][blk] = nfix;
            }
            __syncwarp();
            ok = true;
            for (int blk = 0; blk < kUpBlocks; ++blk) {
                const int nfix = sm.res[warp][blk];
                if (nfix < 0 || nfix > 10) { ok = false; break; }
                rs += nfix;
            }
        }
        if (!ok) continue;                                    // warp-uniform
        uint32_t slot = 0;
        if (lane == 0) slot = atomicAdd(&ctr->n_rec, 1u);
        slot = __shfl_sync(0xFFFFFFFFu, slot, 0);
        if (slot >= rec_cap) { if (lane == 0) ctr->overflow = 1u; continue; }
        RecC &r = rec[slot];
        for (int w = 0; w < 3; ++w) {
            const int64_t x = int64_t(cd.x) - kSigmaLead + 32 * w + lane;
            const uint32_t bits = __ballot_sync(0xFFFFFFFFu, x >= 0 && phi_diff(phi[x >= 0 ? x : 0], phi[(x >= 0 ? x : 0) + 1]) > 0);
            if (lane == 0) r.sigma[w] = bits;
        }
        if (lane == 0) { r.s = abs_of_index0 + cd.x; r.rs_errors = int16_t(rs); r.uplink = uint8_t(cd.uplink); r.type = uint8_t(type); }
        if (!cd.uplink) {
            const int len = type == 2 ? kLongData : kShortData;
            for (int i = lane; i < len; i += 32) r.data[i] = buf[i];
        } else {
            for (int i = lane; i < kUpData; i += 32) r.data[i] = buf[(i / kUpBlockData) * kUpBlockBytes + i % kUpBlockData];
        }
    }
}

// UAT978.cpp:57 as compiled: memmove(buffer, buffer + done, len - done) -- the count is in BYTES although it is the number
// of SAMPLES left, so only the first half of the unconsumed samples reaches the front and indices [(len-done)/2, len-done)
// keep what they held (oracle/_ref/libref978.so shows it; tests/golden/uat_ref.npz).  Out of place: src -> dst, all 65536.
// One kernel per call prepares the buffer: the previous call's memmove (done / len of that call; len == 0: none pending), then the
// call's new phases behind the `used` samples that are left, and the counters of the call cleared.
__global__ void __launch_bounds__(256) uat_stage_kernel(const uint16_t *src, uint16_t *dst, uint32_t done, uint32_t len, uint32_t used,
                                                        const uint16_t *fresh, uint32_t n_fresh, CountersC *ctr)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) *ctr = CountersC{0u, 0u, 0u, 0u};
    const uint32_t moved_bytes = len - done;
    const uint8_t *sb = reinterpret_cast<const uint8_t *>(src);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < 65536u; i += gridDim.x * blockDim.x) {
        if (i >= used && i < used + n_fresh) { dst[i] = fresh[i - used]; continue; }
        uint32_t lo = sb[2 * i], hi = sb[2 * i + 1];
        if (2 * i < moved_bytes) lo = sb[2 * (done + i)];
        if (2 * i + 1 < moved_bytes) hi = sb[2 * (done + i) + 1];
        dst[i] = uint16_t(lo | (hi << 8));
    }
}

}  // namespace

struct PathC {
    uint16_t *d_lut = nullptr;
    Gf *d_gf = nullptr;
    uint8_t *d_iq = nullptr;      size_t iq_cap = 0;      // upload staging for host feeds
    uint16_t *d_new = nullptr;    size_t new_cap = 0;     // phases of the feed's samples
    uint16_t *d_buf[2] = {nullptr, nullptr};              // UAT978Handler::buffer as the reference sees it (+ slack), double-buffered
    int cur = 0;
    uint16_t *d_phi_out = nullptr; size_t phi_out_cap = 0;
    CandC *d_cand = nullptr;      uint32_t cand_cap = 0;
    RecC *d_rec = nullptr, *h_rec = nullptr; uint32_t rec_cap = 0;
    CountersC *d_ctr = nullptr, *h_ctr = nullptr;
    cudaEvent_t ev[2] = {nullptr, nullptr};
    // UAT978Handler state (UAT978.cpp:106-107)
    uint64_t offset = 0;          // absolute sample of buffer[0]
    uint64_t used = 0;            // samples in the staging buffer
    uint32_t shift_done = 0, shift_len = 0;   // the last call's memmove, applied by the next call's stage kernel (len 0: none)
    std::vector<RecC> pending;    // decoded candidates of the running process_buffer() call, sorted by s
    // timings of the last feed (bench.py)
    float ms_device = 0.f;
    uint32_t calls = 0, launches = 0;
};

void pathc_free(PathC *c)
{
    if (!c) return;
    cudaFree(c->d_lut); cudaFree(c->d_gf); cudaFree(c->d_iq); cudaFree(c->d_new); cudaFree(c->d_buf[0]); cudaFree(c->d_buf[1]); cudaFree(c->d_phi_out);
    cudaFree(c->d_cand); cudaFree(c->d_rec);
    for (auto &e : c->ev) if (e) cudaEventDestroy(e);
    cudaFree(c->d_ctr);
    if (c->h_rec) cudaFreeHost(c->h_rec);
    if (c->h_ctr) cudaFreeHost(c->h_ctr);
    delete c;
}

#define CKC_(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { err = std::string(#call ": ") + cudaGetErrorString(e_); return ADSBCU_ERR_CUDA; } } while (0)

static int pathc_init(PathC *&c, std::string &err)
{
    if (c) return ADSBCU_OK;
    c = new PathC;
    std::vector<uint16_t> lut(65536);
    for (unsigned i = 0; i < 256; ++i) {                   // UAT978.cpp:76-100
        const double dI = i - 127.5;
        for (unsigned q = 0; q < 256; ++q) {
            const double dQ = q - 127.5;
            const double ang = std::atan2(dQ, dI) + 3.14159265358979323846;
            const double scaled = std::round(32768 * ang / 3.14159265358979323846);
            lut[i | (q << 8)] = uint16_t(scaled < 0 ? 0 : scaled > 65535 ? 65535 : scaled);
        }
    }
    Gf gf;
    int sr = 1;
    gf.log[0] = 255; gf.exp[255] = 0;
    for (int i = 0; i < 255; ++i) { gf.log[sr] = uint8_t(i); gf.exp[i] = uint8_t(sr); sr <<= 1; if (sr & 256) sr ^= 0x187; sr &= 255; }
    CKC_(cudaMalloc(&c->d_lut, 65536 * 2));
    CKC_(cudaMemcpy(c->d_lut, lut.data(), 65536 * 2, cudaMemcpyHostToDevice));
    CKC_(cudaMalloc(&c->d_gf, sizeof(Gf)));
    CKC_(cudaMemcpy(c->d_gf, &gf, sizeof(Gf), cudaMemcpyHostToDevice));
    CKC_(cudaMalloc(&c->d_ctr, sizeof(CountersC)));
    CKC_(cudaHostAlloc(&c->h_ctr, sizeof(CountersC), cudaHostAllocDefault));
    for (auto &b : c->d_buf) {
        CKC_(cudaMalloc(&b, (65536 + 512) * sizeof(uint16_t)));
        CKC_(cudaMemset(b, 0, (65536 + 512) * sizeof(uint16_t)));              // std::ranges::fill(buffer, 0) (UAT978.cpp:32)
    }
    c->cand_cap = c->rec_cap = 8192;                                           // per process_buffer() call (<= 65536 samples)
    CKC_(cudaMalloc(&c->d_cand, size_t(c->cand_cap) * sizeof(CandC)));
    CKC_(cudaMalloc(&c->d_rec, size_t(c->rec_cap) * sizeof(RecC)));
    CKC_(cudaHostAlloc(&c->h_rec, size_t(c->rec_cap) * sizeof(RecC), cudaHostAllocDefault));
    for (auto &e : c->ev) CKC_(cudaEventCreate(&e));
    return ADSBCU_OK;
}

void pathc_stats(const PathC *c, float *ms_device, uint32_t *calls, uint32_t *launches)
{
    *ms_device = c ? c->ms_device : 0.f; *calls = c ? c->calls : 0u; *launches = c ? c->launches : 0u;
}

void pathc_reset(PathC *c)
{
    if (!c) return;
    c->offset = 0; c->used = 0; c->shift_done = 0; c->shift_len = 0; c->pending.clear();
    for (auto &b : c->d_buf) if (b) cudaMemset(b, 0, (65536 + 512) * sizeof(uint16_t));
}

namespace {

inline bool fuzzy(uint64_t word, uint64_t expected) { return __builtin_popcountll(word ^ expected) <= 4; }

// 36-bit sync register as the reference's scan would hold it at scan bit `b` of a process_buffer() call on
// the buffer starting at absolute sample B: the last (up to 36) bits shifted in since `fresh_from` are true
// stream bits, older ones are whatever the register held when the scan jumped over a decoded frame.
struct Regs { uint64_t stale0 = 0, stale1 = 0; int64_t fresh_from = 0; };

inline uint32_t sigma_at(const RecC &r, int64_t x)           // x: absolute sample; must lie in the record's window
{
    const int64_t i = x - (int64_t(r.s) - kSigmaLead);
    return (r.sigma[i >> 5] >> (i & 31)) & 1u;
}

}  // namespace

int pathc_feed(PathC *&c, int sm_count, cudaStream_t st, const uint8_t *iq, size_t nbytes, bool on_device, adsbcu_uat_frame *out, size_t cap,
               size_t *n_out, std::string &err)
{
    if (n_out) *n_out = 0;
    const int rc0 = pathc_init(c, err);
    if (rc0 != ADSBCU_OK) return rc0;
    const size_t n = nbytes / 2;
    if (!n) return ADSBCU_OK;
    if (n > 0x7FFF0000ull) { err = "buffer too large"; return ADSBCU_ERR_ARG; }
    c->ms_device = 0.f; c->calls = 0; c->launches = 0;

    // ---- device: phases of the feed's samples (UAT978.cpp:52), once per feed ------------------------------
    if (!on_device && c->iq_cap < n * 2) { cudaFree(c->d_iq); c->d_iq = nullptr; c->iq_cap = 0; CKC_(cudaMalloc(&c->d_iq, n * 2 + 64)); c->iq_cap = n * 2; }
    if (c->new_cap < n) { cudaFree(c->d_new); c->d_new = nullptr; c->new_cap = 0; CKC_(cudaMalloc(&c->d_new, n * 2 + 64)); c->new_cap = n; }
    CKC_(cudaEventRecord(c->ev[0], st));
    if (!on_device) CKC_(cudaMemcpyAsync(c->d_iq, iq, n * 2, cudaMemcpyHostToDevice, st));
    ADSBCU_LAUNCH(uat_phase_kernel, sm_count * 8, 256, 0, st, on_device ? iq : c->d_iq, uint32_t(n), c->d_lut, c->d_new);
    CKC_(cudaGetLastError());
    ++c->launches;

    // ---- host: replay UAT978Handler::HandleData's staging and process_buffer()'s scan ----------------------
    size_t nout = 0;
    bool overflow = false;
    auto find = [&](uint64_t s, int uplink) -> const RecC * {
        auto it = std::lower_bound(c->pending.begin(), c->pending.end(), s, [](const RecC &r, uint64_t v) { return r.s < v; });
        for (; it != c->pending.end() && it->s == s; ++it)
            if (it->uplink == uplink) return &*it;
        return nullptr;
    };
    auto emit = [&](const RecC &r) {
        if (nout < cap) {
            adsbcu_uat_frame &f = out[nout];
            std::memset(&f, 0, sizeof(f));
            f.offset = r.s; f.rs_errors = r.rs_errors; f.updown = r.uplink ? '+' : '-';
            f.len = uint16_t(r.uplink ? kUpData : (r.type == 2 ? kLongData : kShortData));
            std::memcpy(f.data, r.data, f.len);
        } else {
            overflow = true;
        }
        ++nout;
    };
    // One iteration == one process_buffer() call of the reference (UAT978.cpp:48-59).  The staging buffer lives on the
    // device exactly as the reference's does on the host -- including what its short memmove leaves behind -- and every
    // call gets its own sync search and frame decode over that buffer: the positions a call presents are not a window
    // of the stream.
    size_t fed = 0;                                          // samples of this feed already copied into the buffer
    while (fed < n) {
        const uint64_t i = std::min<uint64_t>(65536, c->used + (n - fed));   // UAT978.cpp:52
        // stage: the last call's (short) memmove, then the new phases, out of place into the other buffer
        ADSBCU_LAUNCH(uat_stage_kernel, 64, 256, 0, st, c->d_buf[c->cur], c->d_buf[c->cur ^ 1], c->shift_done, c->shift_len, uint32_t(c->used),
                      c->d_new + fed, uint32_t(i - c->used), c->d_ctr);
        CKC_(cudaGetLastError());
        c->cur ^= 1;
        c->shift_done = 0; c->shift_len = 0;
        ++c->launches;
        uint16_t *buf = c->d_buf[c->cur];
        fed += i - c->used;
        const int64_t len = int64_t(i), lenbits = len / 2 - (kSyncBits + kUpBits);
        // fewer than 36 scannable bits: the reference's process_buffer would return a negative count and its caller would
        // memmove from in front of the buffer (undefined); it needs a feed that ends within 72 samples of the scan limit.
        // Defined here (and in the oracle) as "nothing consumed".
        if (lenbits < kSyncBits) { c->used = i; continue; }
        // candidates the scan can try: index = 2 * startbit + {0, 1} (+ 1 for the second alignment), startbit <= lenbits - 36
        const uint32_t last = uint32_t(2 * (lenbits - kSyncBits + 1) + 3);
        ADSBCU_LAUNCH(uat_sync_kernel, (last + 255) / 256, 256, 0, st, buf, 0u, last, c->d_cand, c->cand_cap, c->d_ctr);
        ADSBCU_LAUNCH(uat_decode_kernel, 128, kDecWarps * 32, 0, st, buf, c->offset, c->d_cand, c->d_rec, c->rec_cap, c->d_ctr, c->d_gf);
        CKC_(cudaGetLastError());
        constexpr uint32_t kFirst = 32;                       // records fetched together with the counters
        CKC_(cudaMemcpyAsync(c->h_ctr, c->d_ctr, sizeof(CountersC), cudaMemcpyDeviceToHost, st));
        CKC_(cudaMemcpyAsync(c->h_rec, c->d_rec, kFirst * sizeof(RecC), cudaMemcpyDeviceToHost, st));
        CKC_(cudaStreamSynchronize(st));
        c->launches += 2; ++c->calls;
        if (c->h_ctr->overflow || c->h_ctr->n_cand > c->cand_cap) { err = "internal candidate list overflow"; return ADSBCU_ERR_OVERFLOW; }
        const uint32_t nrec = c->h_ctr->n_rec;
        if (nrec > kFirst) {
            CKC_(cudaMemcpyAsync(c->h_rec + kFirst, c->d_rec + kFirst, size_t(nrec - kFirst) * sizeof(RecC), cudaMemcpyDeviceToHost, st));
            CKC_(cudaStreamSynchronize(st));
        }
        std::sort(c->h_rec, c->h_rec + nrec, [](const RecC &a, const RecC &b) { return a.s != b.s ? a.s < b.s : a.uplink < b.uplink; });
        c->pending.assign(c->h_rec, c->h_rec + nrec);
        const uint64_t B = c->offset;
        int64_t bit = 0;                                      // the scan's loop variable
        Regs regs;
        // scan bits worth looking at: those from which a decoded candidate could be tried
        auto lo = std::lower_bound(c->pending.begin(), c->pending.end(), B, [](const RecC &r, uint64_t v) { return r.s < v; });
        std::vector<int64_t> sbs;
        for (auto it = lo; it != c->pending.end() && it->s < B + uint64_t(2 * lenbits + 4); ++it) {
            const int64_t rel = int64_t(it->s - B);
            sbs.push_back(rel >> 1);
            if ((rel >> 1) >= 1) sbs.push_back((rel >> 1) - 1);
        }
        std::sort(sbs.begin(), sbs.end());
        sbs.erase(std::unique(sbs.begin(), sbs.end()), sbs.end());
        for (const int64_t sb : sbs) {
            const int64_t b = sb + kSyncBits - 1;             // scan bit at which startbit == sb
            if (b < bit || b < kSyncBits || b >= lenbits) continue;
            // any decoded candidate near 2*sb carries the sign bits the registers are made of
            const RecC *win = nullptr;
            for (int e = 0; e <= 3 && !win; ++e)
                for (int up = 0; up < 2 && !win; ++up) win = find(B + uint64_t(2 * sb + e), up);
            if (!win) continue;
            uint64_t r0 = 0, r1 = 0;
            const int64_t nfresh = b - regs.fresh_from + 1;
            const int64_t ntrue = std::min<int64_t>(kSyncBits, nfresh);
            for (int64_t k = b - ntrue + 1; k <= b; ++k) {
                r0 = (r0 << 1) | sigma_at(*win, int64_t(B) + 2 * k);
                r1 = (r1 << 1) | sigma_at(*win, int64_t(B) + 2 * k + 1);
            }
            if (ntrue < kSyncBits) {
                const uint64_t mask = (1ull << kSyncBits) - 1;
                r0 |= (regs.stale0 << ntrue) & mask;
                r1 |= (regs.stale1 << ntrue) & mask;
            }
            bool taken = false;
            for (int up = 0; up < 2 && !taken; ++up) {        // the ADS-B sync word is tried first, then the uplink one
                const uint64_t word = up ? kUplinkSync : kAdsbSync;
                const bool m0 = fuzzy(r0, word), m1 = fuzzy(r1, word);
                if (!m0 && !m1) continue;
                const uint64_t index = B + uint64_t(2 * sb + (m0 ? 0 : 1));
                const RecC *f0 = find(index, up), *f1 = find(index + 1, up);
                const int rs0 = f0 ? f0->rs_errors : 9999, rs1 = f1 ? f1->rs_errors : 9999;
                const RecC *pick = (f0 && rs0 <= rs1) ? f0 : (f1 && rs1 <= rs0) ? f1 : nullptr;
                if (!pick) continue;
                emit(*pick);
                const int skip = kSyncBits + (up ? kUpBits : (pick->type == 2 ? kLongBytes * 8 : kShortBytes * 8));
                bit = sb + skip + 1;                          // bit = startbit + skip; continue -> ++bit
                regs.stale0 = r0; regs.stale1 = r1; regs.fresh_from = bit;
                taken = true;
            }
        }
        if (bit < lenbits) bit = lenbits;                     // the loop runs to lenbits unless a skip carried it past
        const int64_t done = (bit - kSyncBits) * 2;
        c->offset += uint64_t(done);
        c->used = uint64_t(len - done);
        c->shift_done = uint32_t(done); c->shift_len = uint32_t(len);      // UAT978.cpp:57, applied when the next call stages its buffer
    }
    CKC_(cudaEventRecord(c->ev[1], st));
    CKC_(cudaStreamSynchronize(st));
    cudaEventElapsedTime(&c->ms_device, c->ev[0], c->ev[1]);
    c->pending.clear();
    if (n_out) *n_out = nout;
    if (overflow) { err = "out[] too small for the decoded frames"; return ADSBCU_ERR_OVERFLOW; }
    return ADSBCU_OK;
}

// stage-level: the phase table mapping of UAT978.cpp:52 on `n` samples
int pathc_phase(PathC *&c, int sm_count, cudaStream_t st, const uint8_t *iq, size_t n, uint16_t *phi_out, std::string &err)
{
    const int rc0 = pathc_init(c, err);
    if (rc0 != ADSBCU_OK) return rc0;
    if (!n) return ADSBCU_OK;
    if (c->iq_cap < n * 2) { cudaFree(c->d_iq); c->d_iq = nullptr; c->iq_cap = 0; CKC_(cudaMalloc(&c->d_iq, n * 2 + 64)); c->iq_cap = n * 2; }
    if (c->phi_out_cap < n) { cudaFree(c->d_phi_out); c->d_phi_out = nullptr; c->phi_out_cap = 0; CKC_(cudaMalloc(&c->d_phi_out, n * 2 + 64)); c->phi_out_cap = n; }
    CKC_(cudaMemcpyAsync(c->d_iq, iq, n * 2, cudaMemcpyHostToDevice, st));
    ADSBCU_LAUNCH(uat_phase_kernel, sm_count * 8, 256, 0, st, c->d_iq, uint32_t(n), c->d_lut, c->d_phi_out);
    CKC_(cudaGetLastError());
    CKC_(cudaMemcpyAsync(phi_out, c->d_phi_out, n * 2, cudaMemcpyDeviceToHost, st));
    CKC_(cudaStreamSynchronize(st));
    return ADSBCU_OK;
}

}  // namespace adsbcu
