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

}  // namespace adsbcu
