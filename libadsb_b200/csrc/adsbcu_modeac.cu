// Mode A/C on one mag_buf: the reference's demodulate2400AC (dump1090/demod_2400.c:539-726), run by
// dump1090's main loop right after demodulate2400 on the same buffer when --modeac is set
// (dump1090.c:902-905).  SURVEY 8f rank 3.
//
//   acmag_kernel     the 326 overlap samples in front of the block -> magnitudes (fifo.c:187-196)
//   acstats_kernel   sum of magnitudes and of squared magnitudes over the block's new samples: the exact
//                    64-bit integer sums of magnitude_power_uc8 (dsp/impl/magnitude_power_uc8.c:48-90)
//   acscan_kernel    every start position f1 in [1, n): F1/F2 framing pulses, clock phase estimate, the 20 bit
//                    decisions against the +-3 dB thresholds, quiet/framing masks (:606-688).  State-free; the
//                    float steps use explicit round-to-nearest intrinsics so that no FMA contraction can change
//                    what the reference's scalar C computes.
//   host pass        the only sequential rule: a decoded reply skips the next 69 start positions (:722).
#include "../../include/adsbcu.h"
#include "adsbcu_device.cuh"

#include <algorithm>
#include <cstdio>
#include <cstring>

namespace adsbcu {

namespace {

__global__ void acmag_kernel(const uint8_t *iq, uint32_t n, uint16_t *out, const DevTables *tab)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) out[k] = uint16_t(mag_iq(tab->sq, iq[2 * k], iq[2 * k + 1]));
}

__global__ void __launch_bounds__(256) acstats_kernel(const uint16_t *mag, uint32_t n, unsigned long long *sums)
{
    unsigned long long level = 0, power = 0;
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        const uint32_t v = mag[k];
        level += v;
        power += uint64_t(v * v);
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        level += __shfl_xor_sync(0xFFFFFFFFu, level, d);
        power += __shfl_xor_sync(0xFFFFFFFFu, power, d);
    }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&sums[0], level); atomicAdd(&sums[1], power); }
}

// m: magnitudes, m[0] = first overlap sample (the reference's mag->data); positions f1 in [1, mlen)
__global__ void __launch_bounds__(256) acscan_kernel(const uint16_t *m, uint32_t mlen, const unsigned long long *sums,
                                                     AcRec *rec, uint32_t cap, uint32_t *count)
{
    // noise level of the block (:550-551), the same doubles in every thread
    const double mean_level = double(sums[0]) / 65536.0 / double(mlen);
    const double mean_power = double(sums[1]) / 65536.0 / 65536.0 / double(mlen);
    const double noise_stddev = sqrt(mean_power - mean_level * mean_level);
    const unsigned noise_level = unsigned((mean_power + noise_stddev) * 65535 + 0.5);

    for (uint32_t f1 = 1 + blockIdx.x * blockDim.x + threadIdx.x; f1 < mlen; f1 += gridDim.x * blockDim.x) {
        const unsigned a0 = m[f1], a1 = m[f1 + 1], a2 = m[f1 + 2];
        if (!(m[f1 - 1] < a0)) continue;                                   // not a rising edge
        if (a2 > a0 || a2 > a1) continue;                                  // quiet part not quiet
        const unsigned f1_level = (a0 + a1) / 2;
        if (noise_level * 2 > f1_level) continue;                          // 6 dB above noise

        const float pa = __fmul_rn(float(a0), float(a0)), pb = __fmul_rn(float(a1), float(a1));
        const float fraction = __fdiv_rn(pb, __fadd_rn(pa, pb));
        const float f25 = __fmul_rn(25.0f, __fadd_rn(float(f1), __fmul_rn(fraction, fraction)));
        const unsigned f1_clock = unsigned(double(f25) + 0.5);

        const unsigned f2_clock = f1_clock + 87 * 14, f2 = f2_clock / 25;
        const unsigned b0 = m[f2], b1 = m[f2 + 1], b2 = m[f2 + 2];
        if (!(m[f2 - 1] < b0)) continue;
        if (b2 > b0 || b2 > b1) continue;
        const unsigned f2_level = (b0 + b1) / 2;
        if (noise_level * 2 > f2_level) continue;

        const unsigned top = f1_level > f2_level ? f1_level : f2_level;
        const float midpoint = __fsqrt_rn(float(noise_level * top));       // unsigned product -> float, as in C
        const unsigned signal_threshold = unsigned(double(midpoint) * 1.41421356237309504880 + 0.5);
        const unsigned noise_threshold = unsigned(double(midpoint) / 1.41421356237309504880 + 0.5);

        unsigned bits = 0, bad = 0, clock = f1_clock;
#pragma unroll 4
        for (int bit = 0; bit < 20; ++bit, clock += 87) {
            const unsigned s = clock / 25, x0 = m[s], x1 = m[s + 1], x2 = m[s + 2];
            bits <<= 1;
            if (x2 >= signal_threshold) bad = 1;                           // noisy quiet period
            if (x0 >= signal_threshold || x1 >= signal_threshold) bits |= 1;
            else if (x0 > noise_threshold && x1 > noise_threshold) bad = 1;   // uncertain bit
        }
        if ((bits & 0x80020u) != 0x80020u || (bits & 0x0101Bu) != 0 || bad) continue;

        const unsigned modeac =
            ((bits & 0x40000) ? 0x0010 : 0) | ((bits & 0x20000) ? 0x1000 : 0) | ((bits & 0x10000) ? 0x0020 : 0) |
            ((bits & 0x08000) ? 0x2000 : 0) | ((bits & 0x04000) ? 0x0040 : 0) | ((bits & 0x02000) ? 0x4000 : 0) |
            ((bits & 0x00800) ? 0x0100 : 0) | ((bits & 0x00400) ? 0x0001 : 0) | ((bits & 0x00200) ? 0x0200 : 0) |
            ((bits & 0x00100) ? 0x0002 : 0) | ((bits & 0x00080) ? 0x0400 : 0) | ((bits & 0x00040) ? 0x0004 : 0) |
            ((bits & 0x00004) ? 0x0080 : 0);
        const uint32_t slot = atomicAdd(count, 1u);
        if (slot < cap) rec[slot] = AcRec{f1, f2_clock, modeac};
    }
}

}  // namespace

cudaError_t launch_modeac(const DevTables *tab, int sm_count, cudaStream_t s, const uint8_t *d_iq_block, uint32_t n,
                          const uint8_t *d_iq_overlap, uint32_t n_overlap, uint16_t *d_mag_base, unsigned long long *d_sums,
                          AcRec *d_rec, uint32_t cap, uint32_t *d_count)
{
    // d_mag_base + kAcLead is m[0]; the block's first new sample m[326] sits on a 16-byte boundary
    uint16_t *m = d_mag_base + kAcLead;
    cudaMemsetAsync(m, 0, kOverlap * sizeof(uint16_t), s);
    cudaMemsetAsync(d_sums, 0, 2 * sizeof(unsigned long long), s);
    cudaMemsetAsync(d_count, 0, sizeof(uint32_t), s);
    cudaMemsetAsync(m + kOverlap + n, 0, 64, s);
    if (n_overlap) ADSBCU_LAUNCH(acmag_kernel, (n_overlap + 255) / 256, 256, 0, s, d_iq_overlap, n_overlap, m + (kOverlap - n_overlap), tab);
    cudaError_t e = launch_magnitude(d_iq_block, n, m + kOverlap, tab, s);
    if (e != cudaSuccess) return e;
    ADSBCU_LAUNCH(acstats_kernel, sm_count * 4, 256, 0, s, m + kOverlap, n, d_sums);
    ADSBCU_LAUNCH(acscan_kernel, sm_count * 8, 256, 0, s, m, n, d_sums, d_rec, cap, d_count);
    return cudaGetLastError();
}

// ---- adsbcu_1090_modeac_block_static ------------------------------------------------------------------------
// The same kernels as above with every piece of scratch (magnitudes, level / power sums, candidate records) in ONE
// file-scope device array, for callers that may not allocate (CUDA graph capture, serving loops).  No counterpart in
// the reference.  The array is sized for the reference's own mag_buf
// (MODES_MAG_BUF_SAMPLES = 131072 new samples, dump1090.h:92-94) and every GPU that loads the library holds it until
// the process exits.  All calls share it: two calls that overlap on the GPU race.
namespace {

constexpr uint32_t kStaticMaxSamples = ADSBCU_MODEAC_STATIC_MAX_SAMPLES;
constexpr uint32_t kStaticRecCap = kStaticMaxSamples / 16;                 // the ratio adsbcu_1090_modeac_block uses
constexpr size_t kStaticHeadBytes = 256;                                   // sums[2] (u64), candidate counter (u32)
constexpr size_t kStaticRecBytes = size_t(kStaticRecCap) * sizeof(AcRec);
constexpr size_t kStaticRankBytes = size_t(kStaticRecCap) * sizeof(uint32_t);   // records in start-position order (acfinish_kernel)
constexpr size_t kStaticMagBytes = (size_t(kAcLead) + kOverlap + kStaticMaxSamples + 64) * sizeof(uint16_t) + 64;
constexpr size_t kStaticBytes = (kStaticHeadBytes + kStaticRecBytes + kStaticRankBytes + kStaticMagBytes + 255) / 256 * 256;   // ~386 KiB
static_assert(kStaticRecBytes % 256 == 0 && kStaticRankBytes % 256 == 0, "the magnitudes start on a 256-byte boundary");
static_assert(kStaticRecCap <= 8192 && kStaticMaxSamples + kOverlap < (1u << 18), "acfinish_kernel packs (position, record) into 31 bits");

static __device__ __align__(256) unsigned char g_ac_scratch[kStaticBytes];

// In host code g_ac_scratch is only a stand-in for the symbol; the device address comes from cudaGetSymbolAddress,
// asked once per device with that device current.
constexpr int kStaticMaxDevices = 64;
unsigned char *g_ac_scratch_dev[kStaticMaxDevices];                        // [device] -> address of g_ac_scratch there

cudaError_t static_scratch(int device, unsigned char **out)
{
    unsigned char *p = __atomic_load_n(&g_ac_scratch_dev[device], __ATOMIC_ACQUIRE);
    if (!p) {
        void *addr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&addr, g_ac_scratch);   // the current device's copy
        if (e != cudaSuccess) return e;
        p = static_cast<unsigned char *>(addr);
        __atomic_store_n(&g_ac_scratch_dev[device], p, __ATOMIC_RELEASE);  // racing first calls store the same value
    }
    *out = p;
    return cudaSuccess;
}

// One CTA: candidate records in start-position order, the reference's only sequential rule (a decoded reply skips the
// next 69 start positions, demod_2400.c:722) and the adsbcu_msg records (decodeModeAMessage, mode_ac.c:123-158) --
// what adsbcu_1090_modeac_block does on the host after its download.
// result[0] = replies found (may exceed cap: only the first cap were written), result[1] = ADSBCU_STATIC_* flags
__global__ void __launch_bounds__(1024) acfinish_kernel(const AcRec *rec, const uint32_t *count, uint32_t rec_cap, uint32_t *ranked,
                                                         unsigned long long block_timestamp, adsbcu_msg *out, uint32_t cap,
                                                         uint32_t *result)
{
    __shared__ uint32_t key[kStaticRecCap];           // f1 of record i; then, by rank: f1 << 13 | i; then: out[] index << 13 | i
    __shared__ uint32_t nmsg_s;
    const uint32_t found = *count;
    const uint32_t n = found > rec_cap ? 0u : found;  // candidate overflow: no replies, flag set (the host path returns an error)
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) key[i] = rec[i].f1;
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {          // f1 is unique: rank = records in front
        const uint32_t k = key[i];
        uint32_t r = 0;
        for (uint32_t j = 0; j < n; ++j) r += key[j] < k;
        ranked[r] = (k << 13) | i;
    }
    __syncthreads();                                  // ranked[] was written by this CTA: visible after the barrier
    for (uint32_t r = threadIdx.x; r < n; r += blockDim.x) key[r] = ranked[r];
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t nmsg = 0, next_ok = 0;
        for (uint32_t r = 0; r < n; ++r) {
            const uint32_t f1 = key[r] >> 13, i = key[r] & 8191u;
            if (f1 < next_ok) { key[r] = ~0u; continue; }
            next_ok = f1 + 20 * 87 / 25 + 1;
            key[r] = (nmsg++ << 13) | i;
        }
        nmsg_s = nmsg;
    }
    __syncthreads();
    for (uint32_t r = threadIdx.x; r < n; r += blockDim.x) {
        const uint32_t k = key[r] >> 13;
        if (key[r] == ~0u || k >= cap) continue;      // dropped by the skip rule, or out[] is full
        const AcRec a = rec[key[r] & 8191u];
        adsbcu_msg m;
        memset(&m, 0, sizeof(m));
        m.timestamp = block_timestamp + a.f2_clock / 5;                // 60 MHz -> 12 MHz, reported at F2 (demod_2400.c:712)
        m.msgtype = 32;
        m.msgbits = 16;
        m.msg[0] = m.verbatim[0] = uint8_t(a.modeac >> 8);
        m.msg[1] = m.verbatim[1] = uint8_t(a.modeac);
        m.addr = (a.modeac & 0x0000FF7Fu) | kNonIcao;
        out[k] = m;
    }
    if (threadIdx.x == 0) {
        result[0] = nmsg_s;
        result[1] = (found > rec_cap ? uint32_t(ADSBCU_STATIC_CANDIDATE_OVERFLOW) : 0u) |
                    (nmsg_s > cap ? uint32_t(ADSBCU_STATIC_OUT_OVERFLOW) : 0u);
    }
}

}  // namespace
}  // namespace adsbcu

extern "C" int adsbcu_1090_modeac_block_static(adsbcu_handle *h, const uint8_t *iq_dev, size_t nsamples, const uint8_t *overlap_iq_dev,
                                               size_t n_overlap, uint64_t block_timestamp, adsbcu_msg *out_dev, size_t cap,
                                               uint32_t *result_dev, void *cuda_stream)
{
    using namespace adsbcu;
    if (!h) return ADSBCU_ERR_ARG;
    const HandleView v = handle_view(h);
    if (v.host_only) { handle_set_error(h, "this handle has no device (cfg.device == -1)"); return ADSBCU_ERR_ARG; }
    if (nsamples > kStaticMaxSamples) {
        handle_set_error(h, "adsbcu_1090_modeac_block_static: nsamples exceeds ADSBCU_MODEAC_STATIC_MAX_SAMPLES (131072), the size of "
                            "the library's static scratch; use adsbcu_1090_modeac_block for larger blocks");
        return ADSBCU_ERR_ARG;
    }
    if (!result_dev || (!iq_dev && nsamples) || (!out_dev && cap) || (n_overlap && !overlap_iq_dev) || n_overlap > size_t(kOverlap) ||
        (reinterpret_cast<uintptr_t>(iq_dev) & 15u) || cap > 0xFFFFFFFFull || v.device < 0 || v.device >= kStaticMaxDevices) {
        handle_set_error(h, "bad argument (null or misaligned device pointer, more than 326 overlap samples, or cap >= 2^32)");
        return ADSBCU_ERR_ARG;
    }
    cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
    auto fail = [&](cudaError_t e, const char *what) {
        char buf[256];
        std::snprintf(buf, sizeof(buf), "%s: %s", what, cudaGetErrorString(e));
        handle_set_error(h, buf);
        return int(ADSBCU_ERR_CUDA);
    };
    cudaError_t e = cudaSetDevice(v.device);
    if (e != cudaSuccess) return fail(e, "cudaSetDevice");
    unsigned char *scratch = nullptr;
    e = static_scratch(v.device, &scratch);
    if (e != cudaSuccess) return fail(e, "cudaGetSymbolAddress(g_ac_scratch)");
    if (nsamples < 2) {                                // the reference's loop body never runs
        e = cudaMemsetAsync(result_dev, 0, 2 * sizeof(uint32_t), s);
        return e == cudaSuccess ? int(ADSBCU_OK) : fail(e, "cudaMemsetAsync");
    }
    // layout of the scratch: [0,256) sums + counter | candidate records | ranks | magnitudes.  launch_modeac clears the sums, the
    // counter, the overlap magnitudes and the padding behind the block; everything else is written before it is read.
    unsigned long long *sums = reinterpret_cast<unsigned long long *>(scratch);
    uint32_t *count = reinterpret_cast<uint32_t *>(sums + 2);
    AcRec *rec = reinterpret_cast<AcRec *>(scratch + kStaticHeadBytes);
    uint32_t *ranked = reinterpret_cast<uint32_t *>(scratch + kStaticHeadBytes + kStaticRecBytes);
    uint16_t *mag = reinterpret_cast<uint16_t *>(scratch + kStaticHeadBytes + kStaticRecBytes + kStaticRankBytes);
    e = launch_modeac(v.tab, v.sm_count, s, iq_dev, uint32_t(nsamples), overlap_iq_dev, uint32_t(n_overlap), mag, sums, rec,
                      kStaticRecCap, count);
    if (e != cudaSuccess) return fail(e, "launch_modeac");
    ADSBCU_LAUNCH(acfinish_kernel, 1, 1024, 0, s, rec, count, kStaticRecCap, ranked, block_timestamp, out_dev, uint32_t(cap), result_dev);
    e = cudaGetLastError();
    return e == cudaSuccess ? int(ADSBCU_OK) : fail(e, "acfinish_kernel");
}
