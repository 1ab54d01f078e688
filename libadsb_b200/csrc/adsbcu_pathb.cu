// Path B: the 1090 MHz demodulator libadsb itself compiles (reference ADSB1090.cpp, antirez 2.0 MS/s).
// One feed == one ADSB1090Handler::HandleData(span) call (ADSB1090.cpp:158-175): the reference keeps no
// sample state between calls and stops scanning 240 samples before the end of the buffer (:772).
//
//   scanb_kernel   fused: uint8 I/Q -> 129x129 LUT magnitude (ADSB1090.cpp:131-142,163-173) -> preamble
//                  test on every position (:782-811) -> for the survivors both decode attempts of
//                  DetectModeS (plain, and after DetectOutOfPhase/ApplyPhaseCorrection, :683-736,817-853),
//                  the mean-delta filter (:870-881), ModesChecksum / FixSingleBitErrors (:277-332) as
//                  CRC-24 syndromes -> records for everything that can be accepted.
//   host pass      the order-dependent rest, verbatim: skip over an accepted message (:931), the 60 s cache
//                  of recently seen ICAO addresses that validates Address/Parity formats (:195-207,396-435)
//                  and retry-only-if-the-first-attempt-failed (:949-957).
#include "../../include/adsbcu.h"
#include "adsbcu_internal.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

namespace adsbcu {

namespace {

constexpr int kBTile = 4096;                 // start positions per tile
constexpr int kBSpan = 240;                  // samples a start position looks ahead (8 us preamble + 112 bits, 2 per us)
constexpr int kBStage = kBTile + kBSpan + 8; // staged magnitudes: 1 lead sample (m[j-1]) + tile + look-ahead
constexpr int kBThreads = 256;
constexpr int kBCandCap = 1024;

struct RecB {                 // 40 bytes
    uint32_t pos;
    uint32_t crc, addr;
    uint8_t  msg[14];
    uint8_t  msgbits, msgtype;
    int8_t   errorbit;
    uint8_t  attempt;         // 0 plain, 1 phase-corrected retry
    uint8_t  definite;        // 1: DF11/17 with valid (or repaired) CRC; 0: Address/Parity format, needs the cache
    uint8_t  pad[7];
};
static_assert(sizeof(RecB) == 40, "RecB layout");

struct CountersB { unsigned int n_rec, overflow; unsigned long long preambles; };

__device__ __forceinline__ int len_by_type(int t) { return (t == 16 || t == 17 || t == 19 || t == 20 || t == 21) ? 112 : 56; }

// Decode attempt on the 224 samples after the preamble.  `corr`: apply ApplyPhaseCorrection's recurrence.
// Returns false if the attempt yields nothing that could be accepted.
__device__ bool attempt_b(const uint16_t *m /* = &mag[j] */, bool corr, const uint32_t *bit_synd, RecB &r, bool &delta_ok)
{
    const uint16_t *d = m + 16;
    uint64_t h0 = 0, h1 = 0;
    uint32_t sum56 = 0, sum112 = 0;
    int prev = 0;
    uint32_t low = d[0];
    const bool err0 = d[0] == d[1];                       // bits[0] = 2 -> errors = 1 (ADSB1090.cpp:840-846)
    for (int k = 0; k < 112; ++k) {
        const uint32_t high = d[2 * k + 1];
        const uint32_t olow = d[2 * k];                   // uncorrected, for the delta filter (m is restored before it, :856)
        const int dl = int(low) - int(high);
        const int ad = dl < 0 ? -dl : dl;
        int bit;
        if (k > 0 && ad < 256) bit = prev;
        else bit = low > high;
        prev = bit;
        if (bit) { if (k < 56) h0 |= 1ull << (55 - k); else h1 |= 1ull << (111 - k); }
        const int od = int(olow) - int(high);
        const uint32_t aod = uint32_t(od < 0 ? -od : od);
        sum112 += aod;
        if (k < 56) sum56 += aod;
        if (k < 111) {
            uint32_t nxt = d[2 * k + 2];
            if (corr) nxt = (low > high) ? uint16_t((nxt * 5u) / 4u) : uint16_t((nxt * 4u) / 5u);
            low = nxt;
        }
    }
    const int msgtype = int(h0 >> 51) & 31;
    const int bits = len_by_type(msgtype);
    const uint32_t delta = (bits == 112 ? sum112 : sum56) / uint32_t(bits / 2);   // / (msglen * 4)
    delta_ok = delta >= 10u * 255u;
    if (!delta_ok || err0) return false;

    // ModesChecksum(msg) ^ crc field == XOR of the single-bit syndromes of every set bit
    uint32_t syn = 0;
    const int off = bits == 112 ? 0 : 56;
    for (int k = 0; k < bits; ++k) {
        const uint32_t b = k < 56 ? uint32_t(h0 >> (55 - k)) & 1u : uint32_t(h1 >> (111 - k)) & 1u;
        if (b) syn ^= bit_synd[k + off];
    }
    int errorbit = -1;
    bool definite = false;
    uint32_t addr;
    if (msgtype == 11 || msgtype == 17) {
        if (syn != 0) {
            for (int k = 0; k < bits; ++k)
                if (bit_synd[k + off] == syn) { errorbit = k; break; }        // FixSingleBitErrors: first (only) position
            if (errorbit < 0) return false;
            if (errorbit < 56) h0 ^= 1ull << (55 - errorbit); else h1 ^= 1ull << (111 - errorbit);
        }
        definite = true;
        addr = uint32_t(h0 >> 24) & 0xFFFFFFu;
    } else if (msgtype == 0 || msgtype == 4 || msgtype == 5 || msgtype == 16 || msgtype == 20 || msgtype == 21 || msgtype == 24) {
        addr = syn;                                                           // BruteForceAp: crc field ^ checksum
    } else {
        return false;
    }
#pragma unroll
    for (int b = 0; b < 14; ++b)
        r.msg[b] = b < 7 ? uint8_t(h0 >> (48 - 8 * b)) : uint8_t(h1 >> (48 - 8 * (b - 7)));
    const int nb = bits / 8;
    r.crc = (uint32_t(r.msg[nb - 3]) << 16) | (uint32_t(r.msg[nb - 2]) << 8) | r.msg[nb - 1];
    r.addr = addr;
    r.msgbits = uint8_t(bits); r.msgtype = uint8_t(msgtype); r.errorbit = int8_t(errorbit);
    r.attempt = corr; r.definite = definite;
    return true;
}

__global__ void __launch_bounds__(kBThreads) scanb_kernel(const uint8_t *iq, uint32_t n, const uint16_t *lut_g, const uint32_t *bit_synd_g,
                                                          RecB *rec, uint32_t rec_cap, CountersB *ctr)
{
    __shared__ uint16_t s_lut[129 * 129 + 3];
    __shared__ uint16_t s_mag[kBStage];
    __shared__ uint16_t s_cand[kBCandCap];
    __shared__ uint32_t s_synd[112];
    __shared__ uint32_t s_ncand;
    const int tid = threadIdx.x;
    for (int i = tid; i < 129 * 129; i += kBThreads) s_lut[i] = lut_g[i];
    if (tid < 112) s_synd[tid] = bit_synd_g[tid];
    const uint32_t npos = n - 2 * 120;                       // j < mlen - FullLength*2
    const uint32_t ntiles = (npos + kBTile - 1) / kBTile;
    unsigned long long my_pre = 0;
    for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        __syncthreads();
        const int64_t base = int64_t(tile) * kBTile;         // position of s_mag[1]
        if (tid == 0) s_ncand = 0;
        for (int i = tid; i < kBStage; i += kBThreads) {
            const int64_t x = base - 1 + i;
            uint32_t v = 0;
            if (x >= 0 && x < n) {
                int a = int(iq[2 * x]) - 127, b = int(iq[2 * x + 1]) - 127;
                a = a < 0 ? -a : a; b = b < 0 ? -b : b;
                v = s_lut[a * 129 + b];
            }
            s_mag[i] = uint16_t(v);
        }
        __syncthreads();
        for (int p = tid; p < kBTile; p += kBThreads) {
            if (base + p >= npos) break;
            const uint16_t *m = &s_mag[1 + p];
            const uint32_t m0 = m[0];
            if (m0 <= m[1] || m[1] >= m[2] || m[2] <= m[3] || m[3] >= m0 || m[4] >= m0 || m[5] >= m0 || m[6] >= m0 ||
                m[7] <= m[8] || m[8] >= m[9] || m[9] <= m[6])
                continue;
            const uint32_t high = (m0 + m[2] + m[7] + m[9]) / 6u;
            if (m[4] >= high || m[5] >= high) continue;
            if (m[11] >= high || m[12] >= high || m[13] >= high || m[14] >= high) continue;
            ++my_pre;
            const uint32_t slot = atomicAdd(&s_ncand, 1u);
            if (slot < kBCandCap) s_cand[slot] = uint16_t(p); else ctr->overflow = 1u;
        }
        __syncthreads();
        const uint32_t nc = min(s_ncand, uint32_t(kBCandCap));
        for (uint32_t c = tid; c < nc; c += kBThreads) {
            const uint32_t p = s_cand[c];
            const uint16_t *m = &s_mag[1 + p];
            RecB r;
            bool delta_ok;
            const bool got = attempt_b(m, false, s_synd, r, delta_ok);
            if (got) {
                r.pos = uint32_t(base) + p;
                const uint32_t slot = atomicAdd(&ctr->n_rec, 1u);
                if (slot < rec_cap) rec[slot] = r; else ctr->overflow = 1u;
            }
            // retry with phase correction unless the first attempt is certainly accepted or failed the delta filter
            if (!delta_ok || (got && r.definite)) continue;
            const bool oop = (base + p) != 0 && (m[3] > m[2] / 3 || m[10] > m[9] / 3 || m[6] > m[7] / 3 || m[-1] > m[1] / 3);
            if (!oop) continue;                                   // the retry would repeat the first attempt exactly
            RecB r2;
            if (attempt_b(m, true, s_synd, r2, delta_ok)) {
                r2.pos = uint32_t(base) + p;
                const uint32_t slot = atomicAdd(&ctr->n_rec, 1u);
                if (slot < rec_cap) rec[slot] = r2; else ctr->overflow = 1u;
            }
        }
    }
    if (my_pre) atomicAdd(&ctr->preambles, my_pre);
}

}  // namespace

struct PathB {
    uint16_t *d_lut = nullptr;
    uint8_t *d_iq = nullptr;
    size_t iq_cap = 0;
    RecB *d_rec = nullptr, *h_rec = nullptr;
    uint32_t rec_cap = 0;
    CountersB *d_ctr = nullptr, *h_ctr = nullptr;
    std::unordered_map<uint32_t, std::chrono::system_clock::time_point> icao;   // ADSB1090.cpp:195-207
    uint64_t preambles = 0;
};

void pathb_free(PathB *b)
{
    if (!b) return;
    cudaFree(b->d_lut); cudaFree(b->d_iq); cudaFree(b->d_rec); cudaFree(b->d_ctr);
    if (b->h_rec) cudaFreeHost(b->h_rec);
    if (b->h_ctr) cudaFreeHost(b->h_ctr);
    delete b;
}

void pathb_reset(PathB *b) { if (b) b->icao.clear(); }

#define CKB(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { err = std::string(#call ": ") + cudaGetErrorString(e_); return ADSBCU_ERR_CUDA; } } while (0)

int pathb_feed(PathB *&b, const DevTables *d_tab, int sm_count, cudaStream_t s, const uint8_t *iq, size_t nbytes, adsbcu_msg_b *out,
               size_t cap, size_t *n_out, std::string &err)
{
    if (n_out) *n_out = 0;
    if (!b) {
        b = new PathB;
        std::vector<uint16_t> lut(129 * 129);
        for (int i = 0; i <= 128; i++)
            for (int q = 0; q <= 128; q++) lut[size_t(i * 129 + q)] = uint16_t(std::round(std::sqrt(double(i * i + q * q)) * 360));
        CKB(cudaMalloc(&b->d_lut, lut.size() * 2));
        CKB(cudaMemcpy(b->d_lut, lut.data(), lut.size() * 2, cudaMemcpyHostToDevice));
        CKB(cudaMalloc(&b->d_ctr, sizeof(CountersB)));
        CKB(cudaHostAlloc(&b->h_ctr, sizeof(CountersB), cudaHostAllocDefault));
    }
    const size_t n = nbytes / 2;
    if (n <= 240) return ADSBCU_OK;                    // nothing can be scanned: the reference runs j < size - 240 (zero iterations at 240, and its bound would wrap below)
    if (n > 0xFFFFFFF0ull) { err = "buffer too large"; return ADSBCU_ERR_ARG; }
    if (b->iq_cap < nbytes) {
        cudaFree(b->d_iq); b->d_iq = nullptr; b->iq_cap = 0;
        CKB(cudaMalloc(&b->d_iq, nbytes + 64));
        b->iq_cap = nbytes;
    }
    const uint32_t want = uint32_t(std::max<size_t>(4096, n / 64));
    if (b->rec_cap < want) {
        cudaFree(b->d_rec); if (b->h_rec) cudaFreeHost(b->h_rec);
        b->d_rec = nullptr; b->h_rec = nullptr; b->rec_cap = 0;
        CKB(cudaMalloc(&b->d_rec, size_t(want) * sizeof(RecB)));
        CKB(cudaHostAlloc(&b->h_rec, size_t(want) * sizeof(RecB), cudaHostAllocDefault));
        b->rec_cap = want;
    }
    CKB(cudaMemcpyAsync(b->d_iq, iq, n * 2, cudaMemcpyHostToDevice, s));
    CKB(cudaMemsetAsync(b->d_ctr, 0, sizeof(CountersB), s));
    const uint32_t ntiles = uint32_t((n - 240 + kBTile - 1) / kBTile);
    const uint32_t grid = std::min<uint32_t>(ntiles, uint32_t(sm_count) * 4u);
    ADSBCU_LAUNCH(scanb_kernel, grid, kBThreads, 0, s, b->d_iq, uint32_t(n), b->d_lut, d_tab->bit_synd, b->d_rec, b->rec_cap, b->d_ctr);
    CKB(cudaGetLastError());
    // the counters and the first records come back together: a live-size buffer (262144 bytes, RTLSDR.hpp:55) rarely holds more,
    // so one HandleData call costs one synchronisation
    constexpr uint32_t kFirstB = 512;
    CKB(cudaMemcpyAsync(b->h_ctr, b->d_ctr, sizeof(CountersB), cudaMemcpyDeviceToHost, s));
    CKB(cudaMemcpyAsync(b->h_rec, b->d_rec, size_t(std::min(kFirstB, b->rec_cap)) * sizeof(RecB), cudaMemcpyDeviceToHost, s));
    CKB(cudaStreamSynchronize(s));
    const CountersB c = *b->h_ctr;
    if (c.overflow || c.n_rec > b->rec_cap) { err = "internal candidate list overflow"; return ADSBCU_ERR_OVERFLOW; }
    b->preambles = c.preambles;
    if (c.n_rec > kFirstB) {
        CKB(cudaMemcpyAsync(b->h_rec + kFirstB, b->d_rec + kFirstB, size_t(c.n_rec - kFirstB) * sizeof(RecB), cudaMemcpyDeviceToHost, s));
        CKB(cudaStreamSynchronize(s));
    }
    RecB *recs = b->h_rec;
    std::sort(recs, recs + c.n_rec, [](const RecB &x, const RecB &y) { return x.pos != y.pos ? x.pos < y.pos : x.attempt < y.attempt; });

    // ---- ordered host pass ------------------------------------------------------------------------
    size_t nmsg = 0;
    bool overflow = false;
    uint64_t next_pos = 0;                              // first position the reference's loop would look at next
    for (size_t i = 0; i < c.n_rec;) {
        size_t e = i + 1;
        while (e < c.n_rec && recs[e].pos == recs[i].pos) ++e;
        if (recs[i].pos >= next_pos) {
            for (size_t k = i; k < e; ++k) {            // plain attempt first, the retry only if it was not accepted
                const RecB &r = recs[k];
                bool good = r.definite;
                if (!good) {
                    auto it = b->icao.find(r.addr);
                    good = it != b->icao.end() && (std::chrono::system_clock::now() - it->second) <= std::chrono::seconds{60};
                }
                if (!good) continue;
                if (nmsg < cap) {
                    adsbcu_msg_b &m = out[nmsg];
                    std::memset(&m, 0, sizeof(m));
                    m.pos = r.pos; m.crc = r.crc; m.addr = r.addr;
                    std::memcpy(m.msg, r.msg, 14);
                    m.msgbits = r.msgbits; m.msgtype = r.msgtype; m.errorbit = r.errorbit; m.corrected = r.attempt;
                } else {
                    overflow = true;
                }
                ++nmsg;
                if (r.definite && r.errorbit == -1) b->icao[r.addr] = std::chrono::system_clock::now();   // :584-593
                next_pos = uint64_t(r.pos) + uint64_t(8 + r.msgbits) * 2 + 1;                             // :931 and the loop's j++
                break;
            }
        }
        i = e;
    }
    if (n_out) *n_out = nmsg;
    if (overflow) { err = "out[] too small for the accepted messages"; return ADSBCU_ERR_OVERFLOW; }
    return ADSBCU_OK;
}

}  // namespace adsbcu
