// Small sm_100a kernels of the 1090 MHz path that are not part of the chunk pipeline
// (front / message / order / resolve / level: adsbcu_front.cu, adsbcu_msgs.cu, adsbcu_finish.cu).
//
//   evaluate_kernel  sharded candidate scans only (adsbcu_1090_revive): one warp per parked record whose address
//                    another shard announced; re-slices it from the I/Q samples (the planes of its chunk are gone)
//                    and writes the fully evaluated record (demod_2400.c:223-347, mode_s.c:520-626).
//   revive_kernel    lays those records out over the segments; or_words_kernel ORs two address sets.
//   carry_kernel     keeps the last 326 magnitudes for the next feed (fifo.c:187-196).
//   magnitude_kernel stage-level I/Q -> magnitude (starch_magnitude_uc8); build_lut_kernel, convert_sc16_kernel.
#include "adsbcu_device.cuh"

namespace adsbcu {

namespace {

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

cudaError_t init_kernels() { return cudaSuccess; }

cudaError_t launch_build_lut(DevTables *tab, cudaStream_t s)
{
    ADSBCU_LAUNCH(build_lut_kernel, 256, 256, 0, s, tab);
    return cudaGetLastError();
}

cudaError_t launch_revive(const ScanArgs &a, cudaStream_t s)
{
    ADSBCU_LAUNCH(revive_kernel, 296, 256, 0, s, a);
    ADSBCU_LAUNCH(revive_counts_kernel, (a.nseg + 255) / 256, 256, 0, s, a);
    ADSBCU_LAUNCH(evaluate_kernel, a.nseg, 256, 0, s, a);
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
