// The layer after the CRC stage: bulk field decode and CPR for libadsb's own handler (Path B),
// and the CPR decoders of dump1090/cpr.c (Path A).
//
//   fields_kernel   one thread per frame: everything DecodeModesMessage fills after the CRC stage
//                   (ADSB1090.cpp:525-672).  32-byte record in, 80-byte record out, no shared state.
//   cpr_a_kernel    one thread per record: decodeCPRairborne / decodeCPRsurface / decodeCPRrelative (cpr.c:162-369)
//   cpr_kernel      one thread per even/odd pair: DecodeCpr (ADSB1090.cpp:1079-1121) with the NL table of
//                   :993-1055, in IEEE doubles with explicit round-to-nearest intrinsics so that the compiler
//                   cannot contract a*b+c into an FMA the reference's x86-64 build does not have.
//
// The one libm call on the path, atan2 in the velocity heading (:649-652), is not reproducible bit for bit with
// the device's atan2, and the result is truncated to whole degrees.  The heading therefore comes from a table
// over all 2^22 (ew_dir, ns_dir, ew_velocity, ns_velocity) combinations, filled once per handle by the host's
// own libm, the one the reference links: exact by construction, 8 MiB of HBM.
#include "../../include/adsbcu.h"
#include "adsbcu_internal.h"

#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

namespace adsbcu {

struct Fields {
    void *d_in = nullptr, *d_out = nullptr;
    size_t cap_in = 0, cap_out = 0;         // bytes
    int16_t *d_heading = nullptr;           // [ew_dir][ns_dir][ew_velocity][ns_velocity]
};

namespace {

constexpr uint32_t kHeadingEntries = 1u << 22;

__device__ __forceinline__ int32_t to_i32(double v)
{
    // static_cast<int32_t>(double) as x86-64 executes it: cvttsd2si, out of range -> INT32_MIN
    return (v >= 2147483648.0 || v <= -2147483649.0 || v != v) ? INT32_MIN : __double2int_rz(v);
}

__global__ void __launch_bounds__(256) fields_kernel(const uint4 *in, uint32_t n, uint4 *out, const int16_t *heading_lut)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    // adsbcu_msg_b: pos, crc, addr, msg[14], msgbits, msgtype, errorbit, corrected, reserved
    const uint4 lo = in[2 * k], hi = in[2 * k + 1];
    uint8_t m[11];
    {
        const uint32_t w[3] = {lo.w, hi.x, hi.y};
#pragma unroll
        for (int j = 0; j < 11; ++j) m[j] = uint8_t(w[j >> 2] >> (8 * (j & 3)));
    }
    const int df = hi.z >> 24;              // hi.z = msg[12], msg[13], msgbits, msgtype: the DF as sliced

    int32_t f[20];
#pragma unroll
    for (int j = 0; j < 20; ++j) f[j] = 0;
    enum { CA, METYPE, MESUB, HEADING, FFLAG, RAWLAT, RAWLON, FLIGHT0, FLIGHT1, EWDIR, EWVEL, NSDIR, NSVEL, VELOCITY, FS, DR, UM,
           IDENTITY, ALTITUDE, UNIT };
    f[CA] = f[FS] = m[0] & 7;
    f[METYPE] = m[4] >> 3;
    f[MESUB] = m[4] & 7;
    f[DR] = m[1] >> 3;
    f[UM] = ((m[1] & 7) << 3) | (m[2] >> 5);
    {
        const int a = ((m[3] & 0x80) >> 5) | (m[2] & 0x02) | ((m[2] & 0x08) >> 3);
        const int b = ((m[3] & 0x02) << 1) | ((m[3] & 0x08) >> 2) | ((m[3] & 0x20) >> 5);
        const int c = ((m[2] & 0x01) << 2) | ((m[2] & 0x04) >> 1) | ((m[2] & 0x10) >> 4);
        const int d = ((m[3] & 0x01) << 2) | ((m[3] & 0x04) >> 1) | ((m[3] & 0x10) >> 4);
        f[IDENTITY] = a * 1000 + b * 100 + c * 10 + d;
    }
    if (df == 0 || df == 4 || df == 16 || df == 20) {       // 13-bit altitude code (:440-466)
        if (m[3] & 0x40) f[UNIT] = 1;
        else if (m[3] & 0x10)
            f[ALTITUDE] = ((((m[2] & 31) << 6) | ((m[3] & 0x80) >> 2) | ((m[3] & 0x20) >> 1) | (m[3] & 15)) * 25) - 1000;
    }
    if (df == 17) {
        const int me = f[METYPE], sub = f[MESUB];
        if (me >= 1 && me <= 4) {
            // eight 6-bit characters in msg[5..10]; the reference's charset: 1..26 letters, 32 space, 48..57 digits
            uint64_t v = 0;
#pragma unroll
            for (int j = 5; j <= 10; ++j) v = (v << 8) | m[j];
            uint32_t pack[2] = {0, 0};
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const uint32_t c = uint32_t(v >> (42 - 6 * j)) & 63;
                const uint32_t ch = (c >= 1 && c <= 26) ? c + 64 : (c == 32 || (c >= 48 && c <= 57)) ? c : uint32_t('?');
                pack[j >> 2] |= ch << (8 * (j & 3));
            }
            f[FLIGHT0] = int32_t(pack[0]);
            f[FLIGHT1] = int32_t(pack[1]);
        } else if (me >= 9 && me <= 18) {
            f[FFLAG] = m[6] & 4;
            if (m[5] & 1) f[ALTITUDE] = ((((m[5] >> 1) << 4) | (m[6] >> 4)) * 25) - 1000;   // 12-bit code (:470-486)
            f[RAWLAT] = ((m[6] & 3) << 15) | (m[7] << 7) | (m[8] >> 1);
            f[RAWLON] = ((m[8] & 1) << 16) | (m[9] << 8) | m[10];
        } else if (me == 19 && sub >= 1 && sub <= 4) {
            if (sub <= 2) {
                const int ewd = (m[5] >> 2) & 1, ewv = ((m[5] & 3) << 8) | m[6];
                const int nsd = m[7] >> 7, nsv = ((m[7] & 0x7f) << 3) | (m[8] >> 5);
                f[EWDIR] = ewd; f[EWVEL] = ewv; f[NSDIR] = nsd; f[NSVEL] = nsv;
                f[VELOCITY] = __double2int_rz(__dsqrt_rn(double(nsv * nsv + ewv * ewv)));
                if (f[VELOCITY]) f[HEADING] = heading_lut[(ewd << 21) | (nsd << 20) | (ewv << 10) | nsv];
            } else {
                f[HEADING] = __double2int_rz(__dmul_rn(360.0 / 128, double(((m[5] & 3) << 5) | (m[6] >> 3))));
            }
        }
    }
    uint4 *o = out + 5 * size_t(k);
#pragma unroll
    for (int j = 0; j < 5; ++j) o[j] = make_uint4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
}

// NL(lat): the 58 zone edges of 1090-WP-9-14 as the reference lists them (ADSB1090.cpp:993-1055)
__constant__ double c_nl_edge[58] = {
    10.47047130, 14.82817437, 18.18626357, 21.02939493, 23.54504487, 25.82924707, 27.93898710, 29.91135686, 31.77209708,
    33.53993436, 35.22899598, 36.85025108, 38.41241892, 39.92256684, 41.38651832, 42.80914012, 44.19454951, 45.54626723,
    46.86733252, 48.16039128, 49.42776439, 50.67150166, 51.89342469, 53.09516153, 54.27817472, 55.44378444, 56.59318756,
    57.72747354, 58.84763776, 59.95459277, 61.04917774, 62.13216659, 63.20427479, 64.26616523, 65.31845310, 66.36171008,
    67.39646774, 68.42322022, 69.44242631, 70.45451075, 71.45986473, 72.45884545, 73.45177442, 74.43893416, 75.42056257,
    76.39684391, 77.36789461, 78.33374083, 79.29428225, 80.24923213, 81.19801349, 82.13956981, 83.07199445, 83.99173563,
    84.89166191, 85.75541621, 86.53536998, 87.00000000};

__device__ __forceinline__ int cpr_nl(double lat)
{
    lat = fabs(lat);
    // number of edges <= lat, by bisection over the sorted table (58 entries, 6 probes)
    int lo = 0, hi = 58;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (lat < c_nl_edge[mid]) hi = mid; else lo = mid + 1;
    }
    return 59 - lo;
}
__device__ __forceinline__ int cpr_mod(int a, int b) { const int r = a % b; return r < 0 ? r + b : r; }

__global__ void __launch_bounds__(256) cpr_kernel(const adsbcu_cpr_pair *in, uint32_t n, adsbcu_cpr_fix *out)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const adsbcu_cpr_pair p = in[k];
    const double lat0 = p.even_lat, lon0 = p.even_lon, lat1 = p.odd_lat, lon1 = p.odd_lon;
    adsbcu_cpr_fix r{0, 0, 0};
    // j = floor((59*lat0 - 60*lat1)/131072 + 0.5)
    const int j = __double2int_rd(__dadd_rn(__ddiv_rn(__dsub_rn(__dmul_rn(59.0, lat0), __dmul_rn(60.0, lat1)), 131072.0), 0.5));
    double rlat0 = __dmul_rn(360.0 / 60, __dadd_rn(double(cpr_mod(j, 60)), __ddiv_rn(lat0, 131072.0)));
    double rlat1 = __dmul_rn(360.0 / 59, __dadd_rn(double(cpr_mod(j, 59)), __ddiv_rn(lat1, 131072.0)));
    if (rlat0 >= 270) rlat0 = __dsub_rn(rlat0, 360.0);
    if (rlat1 >= 270) rlat1 = __dsub_rn(rlat1, 360.0);
    const int nl0 = cpr_nl(rlat0), nl1 = cpr_nl(rlat1);
    if (nl0 == nl1) {
        const bool even = p.even_is_newer != 0;
        const double rlat = even ? rlat0 : rlat1;
        const int nl = even ? nl0 : nl1, ni = max(nl - (even ? 0 : 1), 1);
        const int m = __double2int_rd(__dadd_rn(
            __ddiv_rn(__dsub_rn(__dmul_rn(lon0, double(nl - 1)), __dmul_rn(lon1, double(nl))), 131072.0), 0.5));
        double lon = __dmul_rn(__dmul_rn(__ddiv_rn(360.0, double(ni)),
                                         __dadd_rn(double(cpr_mod(m, ni)), __ddiv_rn(even ? lon0 : lon1, 131072.0))),
                               10000000.0);
        if (lon > 1800000000.0) lon = __dsub_rn(lon, 3600000000.0);
        r.lat1e7 = to_i32(__dmul_rn(rlat, 10000000.0));
        r.lon1e7 = to_i32(lon);
        r.valid = 1;
    }
    out[k] = r;
}

// ---- Path A: dump1090/cpr.c.  Every double step goes through an explicit round-to-nearest intrinsic in the
// reference's operation order; fmod and floor are exact on both sides. ------------------------------------
__device__ __forceinline__ double dadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double dsub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double dmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double ddiv(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ int cpr_zones(double lat, int odd) { return max(cpr_nl(lat) - odd, 1); }     // cpr.c:142-146
__device__ __forceinline__ double pfmod(double a, double b) { const double r = fmod(a, b); return r < 0 ? dadd(r, b) : r; }

__device__ __forceinline__ double surface_quadrant(double rlat, double reflat)     // cpr.c:262-278
{
    if (rlat == 0) return reflat < -45 ? -90.0 : reflat > 45 ? 90.0 : rlat;
    return dsub(rlat, reflat) > 45 ? dsub(rlat, 90.0) : rlat;
}

__device__ int cpr_global(const adsbcu_cpr_in &in, bool surface, double &out_lat, double &out_lon)   // cpr.c:162-311
{
    const double span = surface ? 90.0 : 360.0;
    const double lat0 = in.even_lat, lat1 = in.odd_lat, lon0 = in.even_lon, lon1 = in.odd_lon;
    const int j = __double2int_rd(dadd(ddiv(dsub(dmul(59.0, lat0), dmul(60.0, lat1)), 131072.0), 0.5));
    double rlat0 = dmul(ddiv(span, 60.0), dadd(double(cpr_mod(j, 60)), ddiv(lat0, 131072.0)));
    double rlat1 = dmul(ddiv(span, 59.0), dadd(double(cpr_mod(j, 59)), ddiv(lat1, 131072.0)));
    if (surface) {
        rlat0 = surface_quadrant(rlat0, in.reflat);
        rlat1 = surface_quadrant(rlat1, in.reflat);
    } else {
        if (rlat0 >= 270) rlat0 = dsub(rlat0, 360.0);
        if (rlat1 >= 270) rlat1 = dsub(rlat1, 360.0);
    }
    if (rlat0 < -90 || rlat0 > 90 || rlat1 < -90 || rlat1 > 90) return -2;
    const int nl0 = cpr_nl(rlat0), nl1 = cpr_nl(rlat1);
    if (nl0 != nl1) return -1;
    const bool odd = in.fflag != 0;
    const double rlat = odd ? rlat1 : rlat0;
    const int ni = max(nl0 - (odd ? 1 : 0), 1);
    const int m = __double2int_rd(dadd(ddiv(dsub(dmul(lon0, double(nl0 - 1)), dmul(lon1, double(nl0))), 131072.0), 0.5));
    double rlon = dmul(ddiv(span, double(ni)), dadd(double(cpr_mod(m, ni)), ddiv(odd ? lon1 : lon0, 131072.0)));
    if (surface) rlon = dadd(rlon, dmul(floor(ddiv(dadd(dsub(in.reflon, rlon), 45.0), 90.0)), 90.0));
    rlon = dsub(rlon, dmul(floor(ddiv(dadd(rlon, 180.0), 360.0)), 360.0));
    out_lat = rlat;
    out_lon = rlon;
    return 0;
}

__device__ int cpr_relative(const adsbcu_cpr_in &in, bool surface, double &out_lat, double &out_lon)   // cpr.c:325-369
{
    const bool odd = in.fflag != 0;
    const double flat = ddiv(double(in.even_lat), 131072.0), flon = ddiv(double(in.even_lon), 131072.0);
    const double dlat = ddiv(surface ? 90.0 : 360.0, odd ? 59.0 : 60.0);
    const int j = __double2int_rz(dadd(floor(ddiv(in.reflat, dlat)), floor(dsub(dadd(0.5, ddiv(pfmod(in.reflat, dlat), dlat)), flat))));
    double rlat = dmul(dlat, dadd(double(j), flat));
    if (rlat >= 270) rlat = dsub(rlat, 360.0);
    if (rlat < -90 || rlat > 90) return -1;
    if (fabs(dsub(rlat, in.reflat)) > ddiv(dlat, 2.0)) return -1;
    const double dlon = ddiv(surface ? 90.0 : 360.0, double(cpr_zones(rlat, odd ? 1 : 0)));
    const int m = __double2int_rz(dadd(floor(ddiv(in.reflon, dlon)), floor(dsub(dadd(0.5, ddiv(pfmod(in.reflon, dlon), dlon)), flon))));
    double rlon = dmul(dlon, dadd(double(m), flon));
    if (rlon > 180) rlon = dsub(rlon, 360.0);
    if (fabs(dsub(rlon, in.reflon)) > ddiv(dlon, 2.0)) return -1;
    out_lat = rlat;
    out_lon = rlon;
    return 0;
}

__global__ void __launch_bounds__(256) cpr_a_kernel(const adsbcu_cpr_in *in, uint32_t n, adsbcu_cpr_out *out)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const adsbcu_cpr_in r = in[k];
    adsbcu_cpr_out o{0.0, 0.0, 0, 0};
    o.result = r.mode < ADSBCU_CPR_RELATIVE_AIRBORNE ? cpr_global(r, r.mode == ADSBCU_CPR_SURFACE, o.lat, o.lon)
                                                     : cpr_relative(r, r.mode == ADSBCU_CPR_RELATIVE_SURFACE, o.lat, o.lon);
    out[k] = o;
}

bool grow(void *&p, size_t &cap, size_t need)
{
    if (need <= cap) return true;
    cudaFree(p);
    p = nullptr;
    cap = 0;
    const size_t want = need + need / 2;
    if (cudaMalloc(&p, want) != cudaSuccess) return false;
    cap = want;
    return true;
}

int fail(std::string &err, const char *what, cudaError_t e)
{
    err = std::string(what) + ": " + cudaGetErrorString(e);
    return ADSBCU_ERR_CUDA;
}

}  // namespace

int fields_decode(Fields *&f, cudaStream_t s, const adsbcu_msg_b *msgs, size_t n, adsbcu_fields_b *out, std::string &err)
{
    static_assert(sizeof(adsbcu_msg_b) == 32 && sizeof(adsbcu_fields_b) == 80, "record layout");
    if (!f) f = new Fields();
    if (!n) return ADSBCU_OK;
    cudaError_t e;
    if (!f->d_heading) {
        // heading of every (ew_dir, ns_dir, ew_velocity, ns_velocity): ADSB1090.cpp:640-656 on the host's libm
        // (four host threads over the eastbound half; atan2 is odd in its first argument and the conversion
        // truncates towards zero, so the westbound half is 360 - h, or 0 where h is 0)
        std::vector<int16_t> lut(kHeadingEntries);
        auto fill = [&lut](uint32_t begin, uint32_t end) {
            for (uint32_t i = begin; i < end; ++i) {
                const int ewv = int((i >> 10) & 1023);
                int nsv = int(i & 1023);
                if ((i >> 20) & 1) nsv = -nsv;
                const int h = int(std::atan2(double(ewv), double(nsv)) * 360 / (M_PI * 2));
                lut[i] = int16_t(h);
                lut[i | (1u << 21)] = int16_t(h ? 360 - h : 0);
            }
        };
        {
            constexpr uint32_t kHalf = kHeadingEntries / 2, kThreads = 4;
            std::vector<std::thread> pool;
            for (uint32_t t = 1; t < kThreads; ++t) pool.emplace_back(fill, t * (kHalf / kThreads), (t + 1) * (kHalf / kThreads));
            fill(0, kHalf / kThreads);
            for (auto &th : pool) th.join();
        }
        if ((e = cudaMalloc(&f->d_heading, kHeadingEntries * sizeof(int16_t))) != cudaSuccess) return fail(err, "cudaMalloc(heading)", e);
        if ((e = cudaMemcpy(f->d_heading, lut.data(), kHeadingEntries * sizeof(int16_t), cudaMemcpyHostToDevice)) != cudaSuccess)
            return fail(err, "cudaMemcpy(heading)", e);
    }
    if (!grow(f->d_in, f->cap_in, n * sizeof(adsbcu_msg_b)) || !grow(f->d_out, f->cap_out, n * sizeof(adsbcu_fields_b)))
        return fail(err, "cudaMalloc(fields)", cudaGetLastError());
    if ((e = cudaMemcpyAsync(f->d_in, msgs, n * sizeof(adsbcu_msg_b), cudaMemcpyHostToDevice, s)) != cudaSuccess)
        return fail(err, "H2D(fields)", e);
    ADSBCU_LAUNCH(fields_kernel, unsigned((n + 255) / 256), 256, 0, s, static_cast<const uint4 *>(f->d_in), uint32_t(n),
                                                            static_cast<uint4 *>(f->d_out), f->d_heading);
    if ((e = cudaGetLastError()) != cudaSuccess) return fail(err, "fields_kernel", e);
    if ((e = cudaMemcpyAsync(out, f->d_out, n * sizeof(adsbcu_fields_b), cudaMemcpyDeviceToHost, s)) != cudaSuccess)
        return fail(err, "D2H(fields)", e);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(err, "fields sync", e);
    return ADSBCU_OK;
}

int cpr_decode(Fields *&f, cudaStream_t s, const adsbcu_cpr_pair *pairs, size_t n, adsbcu_cpr_fix *out, std::string &err)
{
    if (!f) f = new Fields();
    if (!n) return ADSBCU_OK;
    cudaError_t e;
    if (!grow(f->d_in, f->cap_in, n * sizeof(adsbcu_cpr_pair)) || !grow(f->d_out, f->cap_out, n * sizeof(adsbcu_cpr_fix)))
        return fail(err, "cudaMalloc(cpr)", cudaGetLastError());
    if ((e = cudaMemcpyAsync(f->d_in, pairs, n * sizeof(adsbcu_cpr_pair), cudaMemcpyHostToDevice, s)) != cudaSuccess)
        return fail(err, "H2D(cpr)", e);
    ADSBCU_LAUNCH(cpr_kernel, unsigned((n + 255) / 256), 256, 0, s, static_cast<const adsbcu_cpr_pair *>(f->d_in), uint32_t(n),
                                                         static_cast<adsbcu_cpr_fix *>(f->d_out));
    if ((e = cudaGetLastError()) != cudaSuccess) return fail(err, "cpr_kernel", e);
    if ((e = cudaMemcpyAsync(out, f->d_out, n * sizeof(adsbcu_cpr_fix), cudaMemcpyDeviceToHost, s)) != cudaSuccess)
        return fail(err, "D2H(cpr)", e);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(err, "cpr sync", e);
    return ADSBCU_OK;
}

int cpr_a_decode(Fields *&f, cudaStream_t s, const adsbcu_cpr_in *in, size_t n, adsbcu_cpr_out *out, std::string &err)
{
    static_assert(sizeof(adsbcu_cpr_in) == 40 && sizeof(adsbcu_cpr_out) == 24, "record layout");
    if (!f) f = new Fields();
    if (!n) return ADSBCU_OK;
    for (size_t k = 0; k < n; ++k)
        if (in[k].mode < 0 || in[k].mode > ADSBCU_CPR_RELATIVE_SURFACE) { err = "bad argument (unknown CPR mode)"; return ADSBCU_ERR_ARG; }
    cudaError_t e;
    if (!grow(f->d_in, f->cap_in, n * sizeof(adsbcu_cpr_in)) || !grow(f->d_out, f->cap_out, n * sizeof(adsbcu_cpr_out)))
        return fail(err, "cudaMalloc(cpr)", cudaGetLastError());
    if ((e = cudaMemcpyAsync(f->d_in, in, n * sizeof(adsbcu_cpr_in), cudaMemcpyHostToDevice, s)) != cudaSuccess)
        return fail(err, "H2D(cpr)", e);
    ADSBCU_LAUNCH(cpr_a_kernel, unsigned((n + 255) / 256), 256, 0, s, static_cast<const adsbcu_cpr_in *>(f->d_in), uint32_t(n),
                                                           static_cast<adsbcu_cpr_out *>(f->d_out));
    if ((e = cudaGetLastError()) != cudaSuccess) return fail(err, "cpr_a_kernel", e);
    if ((e = cudaMemcpyAsync(out, f->d_out, n * sizeof(adsbcu_cpr_out), cudaMemcpyDeviceToHost, s)) != cudaSuccess)
        return fail(err, "D2H(cpr)", e);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(err, "cpr sync", e);
    return ADSBCU_OK;
}

// ---- Path A: surveillance fields of decodeModesMessage (mode_s.c:626-779) ----------------------------------------------
namespace {

__device__ __forceinline__ uint32_t msg_bits(const uint8_t *m, int first, int last)     // getbits (mode_s.h:104-150): bits numbered from 1
{
    uint32_t v = 0;
    for (int b = first; b <= last; ++b) v = (v << 1) | ((m[(b - 1) >> 3] >> (7 - ((b - 1) & 7))) & 1u);
    return v;
}

// decodeID13Field (mode_s.c:96-116): the 13-bit identity field as the hex-coded Gillham word 0ABC0...
__device__ __forceinline__ uint32_t id13_to_gillham(uint32_t f)
{
    uint32_t g = 0;
    if (f & 0x1000) g |= 0x0010;   // C1
    if (f & 0x0800) g |= 0x1000;   // A1
    if (f & 0x0400) g |= 0x0020;   // C2
    if (f & 0x0200) g |= 0x2000;   // A2
    if (f & 0x0100) g |= 0x0040;   // C4
    if (f & 0x0080) g |= 0x4000;   // A4
    if (f & 0x0020) g |= 0x0100;   // B1
    if (f & 0x0010) g |= 0x0001;   // D1 or Q
    if (f & 0x0008) g |= 0x0200;   // B2
    if (f & 0x0004) g |= 0x0002;   // D2
    if (f & 0x0002) g |= 0x0400;   // B4
    if (f & 0x0001) g |= 0x0004;   // D4
    return g;
}

// modeAToModeC (mode_ac.c:61-122): the table lookup of the reference indexes by the twelve A/B/C/D bits only
// (track.h:332-341), i.e. it is internalModeAToModeC of the word with every other bit cleared
__device__ __forceinline__ int mode_a_to_c(uint32_t mode_a)
{
    mode_a &= 0x7777u;
    if ((mode_a & 0xFFFF8889u) != 0 || (mode_a & 0x000000F0u) == 0) return -9999;
    uint32_t one = 0, five = 0;
    if (mode_a & 0x0010) one ^= 7;
    if (mode_a & 0x0020) one ^= 3;
    if (mode_a & 0x0040) one ^= 1;
    if ((one & 5) == 5) one ^= 2;
    if (one > 5) return -9999;
    if (mode_a & 0x0002) five ^= 0x0FF;
    if (mode_a & 0x0004) five ^= 0x07F;
    if (mode_a & 0x1000) five ^= 0x03F;
    if (mode_a & 0x2000) five ^= 0x01F;
    if (mode_a & 0x4000) five ^= 0x00F;
    if (mode_a & 0x0100) five ^= 0x007;
    if (mode_a & 0x0200) five ^= 0x003;
    if (mode_a & 0x0400) five ^= 0x001;
    if (five & 1) one = 6 - one;
    return int(five * 5 + one) - 13;
}

__device__ void fields_a_one(const uint8_t *m, int df, adsbcu_fields_a &f)
{
    memset(&f, 0, sizeof(f));
    f.altitude_baro = 0;
    if (df == 0 || df == 4 || df == 16 || df == 20) {                             // AC (mode_s.c:628-637)
        f.AC = uint16_t(msg_bits(m, 20, 32));
        if (f.AC) {
            const uint32_t ac = f.AC;
            int alt;
            if (!(ac & 0x0040)) {                                                  // decodeAC13Field (mode_s.c:122-150)
                f.altitude_baro_unit = 0;
                if (ac & 0x0010) alt = int(((ac & 0x1F80) >> 2) | ((ac & 0x0020) >> 1) | (ac & 0x000F)) * 25 - 1000;
                else { const int c = mode_a_to_c(id13_to_gillham(ac)); alt = c < -12 ? -9999 : 100 * c; }
            } else {
                f.altitude_baro_unit = 1;
                alt = -9999;
            }
            f.altitude_baro = alt;
            if (alt != -9999) f.altitude_baro_valid = 1;
        }
    }
    if (df == 11 || df == 17) {                                                    // CA (:641-661)
        f.CA = uint8_t(msg_bits(m, 6, 8));
        if (f.CA == 0 || f.CA == 6 || f.CA == 7) f.airground = 3;
        else if (f.CA == 4) f.airground = 1;
        else if (f.CA == 5) f.airground = 2;
    }
    if (df == 0) f.CC = uint8_t(msg_bits(m, 7, 7));
    if (df == 18) f.CF = uint8_t(msg_bits(m, 6, 8));
    if (df == 4 || df == 5 || df == 20 || df == 21) {
        f.DR = uint8_t(msg_bits(m, 9, 13));
        f.FS = uint8_t(msg_bits(m, 6, 8));                                         // FS (:678-712)
        f.alert_valid = 1; f.spi_valid = 1;
        switch (f.FS) {
        case 0: f.airground = 3; break;
        case 1: f.airground = 1; break;
        case 2: f.airground = 3; f.alert = 1; break;
        case 3: f.airground = 1; f.alert = 1; break;
        case 4: f.airground = 3; f.alert = 1; f.spi = 1; break;
        case 5: f.airground = 3; f.spi = 1; break;
        default: f.spi_valid = 0; f.alert_valid = 0; break;
        }
        f.UM = uint8_t(msg_bits(m, 14, 19));
    }
    if (df == 5 || df == 21) {                                                     // ID (:715-723)
        f.ID = uint16_t(msg_bits(m, 20, 32));
        if (f.ID) { f.squawk = uint16_t(id13_to_gillham(f.ID)); f.squawk_valid = 1; }
    }
    if (df == 24) { f.KE = uint8_t(msg_bits(m, 4, 4)); f.ND = uint8_t(msg_bits(m, 5, 8)); }
    if (df == 0 || df == 16) {
        f.RI = uint8_t(msg_bits(m, 14, 17));
        f.SL = uint8_t(msg_bits(m, 9, 11));
        f.VS = uint8_t(msg_bits(m, 6, 6));                                         // VS (:772-779)
        f.airground = f.VS ? 1 : 3;
    }
    f.es_pending = (df == 17 || df == 18) ? 1 : 0;
}

__global__ void __launch_bounds__(256) fields_a_kernel(const adsbcu_msg *msgs, uint32_t n, adsbcu_fields_a *out)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    uint8_t m[14];
    const uint8_t *src = reinterpret_cast<const uint8_t *>(&msgs[k]) + 8;        // adsbcu_msg::msg (packed record: byte loads)
    for (int i = 0; i < 14; ++i) m[i] = src[i];
    const int df = reinterpret_cast<const uint8_t *>(&msgs[k])[8 + 14 + 14 + 2];   // adsbcu_msg::msgtype
    adsbcu_fields_a f;
    fields_a_one(m, df, f);
    out[k] = f;
}

}  // namespace

int fields_a_decode(Fields *&f, cudaStream_t s, const adsbcu_msg *msgs, size_t n, adsbcu_fields_a *out, std::string &err)
{
    static_assert(sizeof(adsbcu_fields_a) == 32 && sizeof(adsbcu_msg) == 56, "record layout");
    if (!f) f = new Fields();
    if (!n) return ADSBCU_OK;
    cudaError_t e;
    if (!grow(f->d_in, f->cap_in, n * sizeof(adsbcu_msg)) || !grow(f->d_out, f->cap_out, n * sizeof(adsbcu_fields_a)))
        return fail(err, "cudaMalloc(fields)", cudaGetLastError());
    if ((e = cudaMemcpyAsync(f->d_in, msgs, n * sizeof(adsbcu_msg), cudaMemcpyHostToDevice, s)) != cudaSuccess) return fail(err, "H2D(fields)", e);
    ADSBCU_LAUNCH(fields_a_kernel, unsigned((n + 255) / 256), 256, 0, s, static_cast<const adsbcu_msg *>(f->d_in), uint32_t(n),
                  static_cast<adsbcu_fields_a *>(f->d_out));
    if ((e = cudaGetLastError()) != cudaSuccess) return fail(err, "fields_a_kernel", e);
    if ((e = cudaMemcpyAsync(out, f->d_out, n * sizeof(adsbcu_fields_a), cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail(err, "D2H(fields)", e);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(err, "fields sync", e);
    return ADSBCU_OK;
}

// ---- Path A: extended-squitter payload (decodeExtendedSquitter and its nine decoders, mode_s.c:798-1551) ----------------
// One thread per message; the 56 ME bits sit in one 64-bit register and every field is a shift and a mask of it.  A CTA
// stages its 256 input records and its 256 output records in shared memory so that both directions are coalesced word
// copies (the records are 56 and 120 bytes: neither is a power of two).
namespace {

constexpr uint32_t kNonIcao = 1u << 24;                                        // MODES_NON_ICAO_ADDRESS, dump1090.h:262
// addrtype_t, heading_type_t, nav_altitude_source_t, nav_modes_t, datasource_t (dump1090.h:133-242)
enum : uint8_t { kAdsbIcao = 0, kAdsbIcaoNt = 1, kAdsrIcao = 2, kTisbIcao = 3, kAdsbOther = 4, kAdsrOther = 5, kTisbTrackfile = 6, kTisbOther = 7,
                 kAddrUnknown = 9 };
enum : uint8_t { kHdgGroundTrack = 1, kHdgTrue = 2, kHdgMagnetic = 3, kHdgMagneticOrTrue = 4, kHdgTrackOrHeading = 5 };
enum : uint8_t { kAltAircraft = 2, kAltMcp = 3, kAltFms = 4 };
enum : uint8_t { kNavAutopilot = 1, kNavVnav = 2, kNavAltHold = 4, kNavApproach = 8, kNavLnav = 16, kNavTcas = 32 };
enum : uint8_t { kSrcTisb = 5, kSrcAdsr = 6, kSrcAdsb = 7 };

struct MeBits {
    uint64_t v;                                                                // ME bit 1 is bit 55
    __device__ __forceinline__ uint32_t operator()(int first, int last) const { return uint32_t(v >> (56 - last)) & ((1u << (last - first + 1)) - 1u); }
    __device__ __forceinline__ uint32_t operator()(int bit) const { return uint32_t(v >> (56 - bit)) & 1u; }
    __device__ __forceinline__ uint64_t wide(int first, int last) const { return (v >> (56 - last)) & ((1ull << (last - first + 1)) - 1ull); }   // fields of 32 bits and more
};

__device__ __forceinline__ void non_icao(adsbcu_fields_es &e)                  // setIMF (mode_s.c:832-854)
{
    e.addr |= kNonIcao;
    if (e.addrtype == kAdsbIcao || e.addrtype == kAdsbIcaoNt) e.addrtype = kAdsbOther;
    else if (e.addrtype == kTisbIcao) e.addrtype = kTisbTrackfile;
    else if (e.addrtype == kAdsrIcao) e.addrtype = kAdsrOther;
}

// decodeMovementFieldV0 / V2 (mode_s.c:184-230): midpoints of the piecewise-linear movement scale, computed in double
// and narrowed once, as the reference's `return <double expression>` from a float function does
__device__ __forceinline__ float movement_knots(uint32_t m, bool v2)
{
    const double half = double(m) + 0.5;
    double kt;
    if (m >= 125) kt = 0;
    else if (m == 124) kt = 180;
    else if (m >= 109) kt = __dadd_rn(100, __dmul_rn(half - 109, 5));
    else if (m >= 94) kt = __dadd_rn(70, __dmul_rn(half - 94, 2));
    else if (m >= 39) kt = __dadd_rn(15, half - 39);
    else if (m >= 13) kt = __dadd_rn(2, __dmul_rn(half - 13, 0.50));
    else if (m >= 9) kt = __dadd_rn(1, __dmul_rn(half - 9, 0.25));
    else if (v2) kt = m >= 3 ? __dadd_rn(0.125, __ddiv_rn(__dmul_rn(half - 3, 0.875), 6)) : m >= 2 ? 0.125 / 2 : 0;
    else kt = m >= 2 ? __dadd_rn(0.125, __dmul_rn(half - 2, 0.125)) : 0;
    return __double2float_rn(kt);
}

__device__ __forceinline__ int ac12_altitude(uint32_t f)                       // decodeAC12Field (mode_s.c:156-178); always feet
{
    if (f & 0x10) return int(((f & 0x0FE0) >> 1) | (f & 0x000F)) * 25 - 1000;
    const int c = mode_a_to_c(id13_to_gillham(((f & 0x0FC0) << 1) | (f & 0x003F)));
    return c < -12 ? -9999 : 100 * c;
}

__device__ void es_decode_one(const uint8_t *m, int df, int corrected, adsbcu_fields_es &e)
{
    memset(&e, 0, sizeof(e));
    if (df != 17 && df != 18) return;
    MeBits me;
    me.v = 0;
    for (int i = 0; i < 7; ++i) me.v = (me.v << 8) | m[4 + i];
    e.es = 1;
    e.source = kSrcAdsb;                                                       // mode_s.c:580-585
    e.reliable = corrected == 0;
    e.addr = (uint32_t(m[1]) << 16) | (uint32_t(m[2]) << 8) | m[3];            // AA, :624-626
    const uint32_t ca_cf = m[0] & 7u;
    e.CA_CF = uint8_t(ca_cf);
    const uint32_t metype = me(1, 5);
    e.metype = uint8_t(metype);
    bool imf_bit_counts = false;                                               // check_imf of decodeExtendedSquitter (:1447-1503)
    if (df == 17) {
        e.airground = ca_cf == 4 ? 1 : ca_cf == 5 ? 2 : (ca_cf == 0 || ca_cf >= 6) ? 3 : 0;     // CA, :636-657
    } else {
        switch (ca_cf) {
        case 0: e.addrtype = kAdsbIcaoNt; break;
        case 1: e.addrtype = kAdsbOther; e.addr |= kNonIcao; break;
        case 2: e.source = kSrcTisb; e.addrtype = kTisbIcao; imf_bit_counts = true; break;
        case 3: e.source = kSrcTisb; e.addrtype = kTisbIcao; if (me(1)) non_icao(e); return;      // coarse TIS-B: nothing else is read
        case 5: e.source = kSrcTisb; e.addrtype = kTisbOther; e.addr |= kNonIcao; break;
        case 6: e.source = kSrcAdsr; e.addrtype = kAdsrIcao; imf_bit_counts = true; break;
        default: e.addrtype = kAddrUnknown; e.addr |= kNonIcao; return;
        }
    }
    uint32_t valid = 0;
    if (metype >= 1 && metype <= 4) {                                          // identification and category (:798-829)
        e.mesub = uint8_t(me(6, 8));
        bool good = true;
        for (int k = 0; k < 8; ++k) {
            const uint32_t six = me(9 + 6 * k, 14 + 6 * k);
            // ais_charset.c:3 - "@A-Z[\]^_" | " !"#$%&'()*+,-./" | "0-9:;<=>?": codes 0..31 are '@' + code, 32..63 are ' ' + (code - 32)
            const char c = char(six < 32 ? 0x40 + six : six);
            e.callsign[k] = c;
            good = good && ((c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9') || c == ' ');
        }
        valid |= ADSBCU_ES_CATEGORY | (good ? ADSBCU_ES_CALLSIGN : 0u);
        e.category = uint8_t(((0x0E - metype) << 4) | e.mesub);
    } else if (metype == 19) {                                                 // airborne velocity (:856-963)
        const uint32_t sub = me(6, 8);
        e.mesub = uint8_t(sub);
        if (sub >= 1 && sub <= 4) {
            if (imf_bit_counts && me(9)) non_icao(e);
            valid |= ADSBCU_ES_NAC_V;
            e.nac_v = uint8_t(me(11, 13));
            if (sub <= 2) {
                const uint32_t ew_raw = me(15, 24), ns_raw = me(26, 35);
                if (ew_raw && ns_raw) {
                    const int scale = sub == 2 ? 4 : 1;
                    const int ew = int(ew_raw - 1) * (me(14) ? -scale : scale), ns = int(ns_raw - 1) * (me(25) ? -scale : scale);
                    // sqrtf((ns*ns) + (ew*ew) + 0.5): the sum is a double, narrowed for sqrtf
                    const float gs = __fsqrt_rn(__double2float_rn(__dadd_rn(double(ns * ns + ew * ew), 0.5)));
                    e.gs_v0 = e.gs_v2 = e.gs_selected = gs;
                    valid |= ADSBCU_ES_GS;
                    if (gs > 0) {
                        // float ground_track = atan2(ew_vel, ns_vel) * 180.0 / M_PI
                        float track = __double2float_rn(__ddiv_rn(__dmul_rn(atan2(double(ew), double(ns)), 180.0), 3.14159265358979323846));
                        if (track < 0) track = __fadd_rn(track, 360.f);
                        e.heading = track;
                        e.heading_type = kHdgGroundTrack;
                        valid |= ADSBCU_ES_HEADING;
                    }
                }
            } else {
                if (me(14)) { valid |= ADSBCU_ES_HEADING; e.heading = float(me(15, 24)) * (360.f / 1024.f); e.heading_type = kHdgMagneticOrTrue; }
                const uint32_t airspeed = me(26, 35);
                if (airspeed) {
                    const uint32_t kt = (airspeed - 1) * (sub == 4 ? 4u : 1u);
                    if (me(25)) { valid |= ADSBCU_ES_TAS; e.tas = uint16_t(kt); } else { valid |= ADSBCU_ES_IAS; e.ias = uint16_t(kt); }
                }
            }
            const uint32_t vr = me(38, 46);
            if (vr) {
                const int fpm = int(vr - 1) * (me(37) ? -64 : 64);
                if (me(36)) { e.baro_rate = fpm; valid |= ADSBCU_ES_BARO_RATE; } else { e.geom_rate = fpm; valid |= ADSBCU_ES_GEOM_RATE; }
            }
            const uint32_t delta = me(50, 56);
            if (delta) { valid |= ADSBCU_ES_GEOM_DELTA; e.geom_delta = int(delta - 1) * (me(49) ? -25 : 25); }
        }
    } else if (metype >= 5 && metype <= 8) {                                   // surface position (:965-1001)
        e.airground = 1;
        valid |= ADSBCU_ES_CPR;
        e.cpr_type = 0;
        const uint32_t movement = me(6, 12);
        if (movement > 0 && movement < 125) {
            valid |= ADSBCU_ES_GS;
            e.gs_selected = e.gs_v0 = movement_knots(movement, false);
            e.gs_v2 = movement_knots(movement, true);
        }
        if (me(13)) { valid |= ADSBCU_ES_HEADING; e.heading = float(me(14, 20)) * (360.f / 128.f); e.heading_type = kHdgTrackOrHeading; }
        if (imf_bit_counts && me(21)) non_icao(e);
        e.cpr_odd = uint8_t(me(22));
        e.cpr_lat = me(23, 39);
        e.cpr_lon = me(40, 56);
    } else if (metype == 0 || (metype >= 9 && metype <= 18) || (metype >= 20 && metype <= 22)) {   // airborne position (:1003-1092)
        const uint32_t ss = me(6, 7);
        if (ss == 1 || ss == 2) { valid |= ADSBCU_ES_ALERT; e.alert = 1; }
        else { valid |= ADSBCU_ES_ALERT | ADSBCU_ES_SPI; e.spi = ss == 3; }
        if (imf_bit_counts) { if (me(8)) non_icao(e); }
        else { valid |= ADSBCU_ES_NIC_B; e.nic_b = uint8_t(me(8)); }
        const uint32_t ac12 = me(9, 20);
        if (metype != 0) {
            e.cpr_lat = me(23, 39);
            e.cpr_lon = me(40, 56);
            const bool stuck = ac12 == 0 && e.cpr_lon == 0 && (e.cpr_lat & 0x0fff) == 0 && metype == 15;   // :1056, not used for positioning
            if (!stuck) { valid |= ADSBCU_ES_CPR; e.cpr_type = 1; e.cpr_odd = uint8_t(me(22)); }
        }
        if (ac12 && e.airground != 1) {
            const int alt = ac12_altitude(ac12);
            if (alt != -9999) {
                if (e.airground == 0) e.airground = 3;
                if (metype >= 20) { e.altitude_geom = alt; valid |= ADSBCU_ES_ALTITUDE_GEOM; }
                else { e.altitude_baro = alt; valid |= ADSBCU_ES_ALTITUDE_BARO; }
            }
        }
    } else if (metype == 23) {                                                 // test message (:1094-1107)
        e.mesub = uint8_t(me(6, 8));
        const uint32_t id = me(9, 21);
        if (e.mesub == 7 && id) { valid |= ADSBCU_ES_SQUAWK; e.squawk = uint16_t(id13_to_gillham(id)); }
    } else if (metype == 28) {                                                 // aircraft status (:1109-1129)
        e.mesub = uint8_t(me(6, 8));
        if (e.mesub == 1) {
            valid |= ADSBCU_ES_EMERGENCY;
            e.emergency = uint8_t(me(9, 11));
            const uint32_t id = me(12, 24);
            if (id) { valid |= ADSBCU_ES_SQUAWK; e.squawk = uint16_t(id13_to_gillham(id)); }
            if (imf_bit_counts && me(56)) non_icao(e);
        }
    } else if (metype == 29) {                                                 // target state and status (:1131-1324)
        const uint32_t sub = me(6, 7);
        e.mesub = uint8_t(sub);
        if (imf_bit_counts && me(51)) non_icao(e);
        if ((sub == 0 && me(11) == 0) || sub == 1) {
            valid |= ADSBCU_ES_NAC_P | ADSBCU_ES_NIC_BARO;                     // bits 40-46 mean the same in both versions
            e.nac_p = uint8_t(me(40, 43));
            e.nic_baro = uint8_t(me(44));
            e.sil = uint8_t(me(45, 46));
            e.sil_type = 1;
        }
        if (sub == 0 && me(11) == 0) {                                         // version 1 layout
            const uint32_t vsrc = me(8, 9), vmode = me(14, 15), hsrc = me(26, 27), hmode = me(38, 39), tcas = me(52, 53);
            const uint8_t src = vsrc == 1 ? kAltMcp : vsrc == 2 ? kAltAircraft : vsrc == 3 ? kAltFms : 0;
            uint8_t modes = 0;
            e.nav_altitude_source = src;
            if (vmode == 1 || vmode == 2) {
                valid |= ADSBCU_ES_NAV_MODES;
                modes |= src == kAltFms ? kNavVnav : (vmode == 2 && src == kAltAircraft) ? kNavAltHold : kNavAutopilot;
            }
            const int alt = -1000 + 100 * int(me(16, 25));
            if (src == kAltMcp) { valid |= ADSBCU_ES_NAV_MCP_ALTITUDE; e.nav_mcp_altitude = alt; }
            else if (src == kAltFms) { valid |= ADSBCU_ES_NAV_FMS_ALTITUDE; e.nav_fms_altitude = alt; }
            if (hsrc) {
                valid |= ADSBCU_ES_NAV_HEADING;
                e.nav_heading = float(me(28, 36));
                e.nav_heading_type = me(37) ? kHdgGroundTrack : kHdgMagneticOrTrue;
            }
            if (hmode == 1 || hmode == 2) { valid |= ADSBCU_ES_NAV_MODES; modes |= hsrc == 3 ? kNavLnav : kNavAutopilot; }
            if (tcas != 0) valid |= ADSBCU_ES_NAV_MODES;                        // 0: TCAS assumed, but it does not make the modes valid
            if (tcas != 1) modes |= kNavTcas;
            e.nav_modes = modes;
            valid |= ADSBCU_ES_EMERGENCY;
            e.emergency = uint8_t(me(54, 56));
        } else if (sub == 1) {                                                 // version 2 layout
            const uint32_t alt_bits = me(10, 20), baro_bits = me(21, 29);
            if (alt_bits) {
                const int alt = int(alt_bits - 1) * 32;
                if (me(9)) { valid |= ADSBCU_ES_NAV_FMS_ALTITUDE; e.nav_fms_altitude = alt; }
                else { valid |= ADSBCU_ES_NAV_MCP_ALTITUDE; e.nav_mcp_altitude = alt; }
            }
            if (baro_bits) { valid |= ADSBCU_ES_NAV_QNH; e.nav_qnh = __double2float_rn(__dadd_rn(800.0, __dmul_rn(double(baro_bits - 1), 0.8))); }
            if (me(30)) { valid |= ADSBCU_ES_NAV_HEADING; e.nav_heading = float(me(31, 39)) * (180.f / 256.f); e.nav_heading_type = kHdgMagneticOrTrue; }
            if (me(47)) {
                valid |= ADSBCU_ES_NAV_MODES;
                e.nav_modes = uint8_t((me(48) ? kNavAutopilot : 0) | (me(49) ? kNavVnav : 0) | (me(50) ? kNavAltHold : 0) | (me(52) ? kNavApproach : 0) |
                                      (me(53) ? kNavTcas : 0) | (me(54) ? kNavLnav : 0));
            }
        }
    } else if (metype == 31) {                                                 // operational status (:1326-1445)
        const uint32_t sub = me(6, 8);
        e.mesub = uint8_t(sub);
        if (imf_bit_counts && me(56)) non_icao(e);
        if (sub <= 1) {
            const uint32_t version = me(41, 43);
            const bool surface = sub == 1;
            uint32_t ops = 0;
            valid |= ADSBCU_ES_OPSTATUS;
            e.ops_version = uint8_t(version);
            if (version == 0) {
                if (!surface && me(9, 10) == 0) ops |= (me(12) ? 0u : ADSBCU_OPS_CC_ACAS) | (me(13) ? ADSBCU_OPS_CC_CDTI : 0u);
            } else if (version <= 2) {
                const bool v2 = version == 2;
                if (me(25, 26) == 0) {                                         // operational mode codes
                    ops |= (me(27) ? ADSBCU_OPS_OM_ACAS_RA : 0u) | (me(28) ? ADSBCU_OPS_OM_IDENT : 0u) | (me(29) ? ADSBCU_OPS_OM_ATC : 0u);
                    if (v2) { ops |= me(30) ? ADSBCU_OPS_OM_SAF : 0u; valid |= ADSBCU_ES_SDA; e.sda = uint8_t(me(31, 32)); }
                }
                if (me(9, 10) == 0 && (v2 || me(13, 14) == 0)) {               // capability class codes
                    if (!surface) {
                        if (v2) ops |= (me(11) ? ADSBCU_OPS_CC_ACAS : 0u) | (me(12) ? ADSBCU_OPS_CC_1090_IN : 0u) | (me(19) ? ADSBCU_OPS_CC_UAT_IN : 0u);
                        else ops |= (me(11) ? 0u : ADSBCU_OPS_CC_ACAS) | (me(12) ? ADSBCU_OPS_CC_CDTI : 0u);
                        ops |= (me(15) ? ADSBCU_OPS_CC_ARV : 0u) | (me(16) ? ADSBCU_OPS_CC_TS : 0u);
                        e.ops_cc_tc = uint8_t(me(17, 18));
                    } else {
                        ops |= (me(11) ? ADSBCU_OPS_CC_POA : 0u) | (me(15) ? ADSBCU_OPS_CC_B2_LOW : 0u);
                        valid |= ADSBCU_ES_CC_LW;
                        e.ops_cc_lw = uint8_t(me(21, 24));
                        if (v2) {
                            ops |= (me(12) ? ADSBCU_OPS_CC_1090_IN : 0u) | (me(16) ? ADSBCU_OPS_CC_UAT_IN : 0u);
                            valid |= ADSBCU_ES_NAC_V | ADSBCU_ES_NIC_C;
                            e.nac_v = uint8_t(me(17, 19));
                            e.nic_c = uint8_t(me(20));
                            e.ops_cc_antenna_offset = uint8_t(me(33, 40));
                        } else {
                            ops |= me(12) ? ADSBCU_OPS_CC_CDTI : 0u;
                        }
                    }
                }
                valid |= ADSBCU_ES_NIC_A | ADSBCU_ES_NAC_P;
                e.nic_a = uint8_t(me(44));
                e.nac_p = uint8_t(me(45, 48));
                e.sil = uint8_t(me(51, 52));
                e.sil_type = v2 ? (me(55) ? 2 : 3) : 1;
                e.ops_hrd = me(54) ? kHdgMagnetic : kHdgTrue;
                if (!surface) {
                    valid |= ADSBCU_ES_NIC_BARO;
                    e.nic_baro = uint8_t(me(53));
                    if (v2) { valid |= ADSBCU_ES_GVA; e.gva = uint8_t(me(49, 50)); }
                } else {
                    e.ops_tah = me(53) ? e.ops_hrd : kHdgGroundTrack;
                }
            }
            e.ops_flags = uint16_t(ops);
        }
    } else if (metype != 24 && metype != 30) {
        e.reliable = 0;                                                        // "Dubious" (:1546-1549)
    }
    e.valid = valid;
}

constexpr int kEsThreads = 256;

__global__ void __launch_bounds__(kEsThreads) fields_es_kernel(const uint32_t *msgs, uint32_t n, uint32_t *out)
{
    constexpr uint32_t kInWords = sizeof(adsbcu_msg) / 4, kOutWords = sizeof(adsbcu_fields_es) / 4;
    __shared__ uint32_t s_in[kEsThreads * kInWords];
    __shared__ uint32_t s_out[kEsThreads * kOutWords];
    const uint32_t first = blockIdx.x * kEsThreads;
    const uint32_t count = min(uint32_t(kEsThreads), n - first);
    for (uint32_t w = threadIdx.x; w < count * kInWords; w += kEsThreads) s_in[w] = msgs[size_t(first) * kInWords + w];
    __syncthreads();
    if (threadIdx.x < count) {
        const uint8_t *rec = reinterpret_cast<const uint8_t *>(s_in + threadIdx.x * kInWords);
        adsbcu_fields_es e;
        es_decode_one(rec + 8, rec[8 + 14 + 14 + 2], rec[8 + 14 + 14 + 1], e);      // adsbcu_msg::msg, ::msgtype, ::correctedbits
        const uint32_t *src = reinterpret_cast<const uint32_t *>(&e);
        for (uint32_t w = 0; w < kOutWords; ++w) s_out[threadIdx.x * kOutWords + w] = src[w];
    }
    __syncthreads();
    for (uint32_t w = threadIdx.x; w < count * kOutWords; w += kEsThreads) out[size_t(first) * kOutWords + w] = s_out[w];
}

}  // namespace

int fields_es_decode(Fields *&f, cudaStream_t s, const adsbcu_msg *msgs, size_t n, adsbcu_fields_es *out, std::string &err)
{
    static_assert(sizeof(adsbcu_fields_es) == 120 && sizeof(adsbcu_msg) == 56, "record layout");
    if (!f) f = new Fields();
    if (!n) return ADSBCU_OK;
    cudaError_t e;
    if (!grow(f->d_in, f->cap_in, n * sizeof(adsbcu_msg)) || !grow(f->d_out, f->cap_out, n * sizeof(adsbcu_fields_es)))
        return fail(err, "cudaMalloc(fields)", cudaGetLastError());
    if ((e = cudaMemcpyAsync(f->d_in, msgs, n * sizeof(adsbcu_msg), cudaMemcpyHostToDevice, s)) != cudaSuccess) return fail(err, "H2D(es)", e);
    ADSBCU_LAUNCH(fields_es_kernel, unsigned((n + kEsThreads - 1) / kEsThreads), kEsThreads, 0, s, static_cast<const uint32_t *>(f->d_in), uint32_t(n),
                  static_cast<uint32_t *>(f->d_out));
    if ((e = cudaGetLastError()) != cudaSuccess) return fail(err, "fields_es_kernel", e);
    if ((e = cudaMemcpyAsync(out, f->d_out, n * sizeof(adsbcu_fields_es), cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail(err, "D2H(es)", e);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(err, "es sync", e);
    return ADSBCU_OK;
}

// ---- Path A: Comm-B payload of DF20 / DF21 (decodeCommB and its register decoders, comm_b.c:50-985) ----------------------
// One thread per reply.  The reference walks ten scoring functions twice (score, then store); here every register's
// plausibility score is computed once from the 56 MB bits in a 64-bit register, the unique maximum picks the register, and
// only that one is decoded into the record.  Records are staged in shared memory like the ES decoder's.
namespace {

enum : uint8_t { kCbUnknown = 0, kCbAmbiguous = 1, kCbNotDecoded = 2, kCbEmpty = 3, kCbDatalinkCaps = 4, kCbGicbCaps = 5, kCbIdent = 6, kCbAcasRa = 7,
                 kCbVerticalIntent = 8, kCbTrackTurn = 9, kCbHeadingSpeed = 10, kCbMrar = 11, kCbAirbornePosition = 12 };   // commb_format_t

// 0 = not this register.  `in range or out` helpers keep the scoring tables readable.
__device__ __forceinline__ bool within(uint32_t v, uint32_t lo, uint32_t hi) { return v >= lo && v <= hi; }

__device__ __forceinline__ int score_caps17(const MeBits &mb)                    // BDS 1,7 (comm_b.c:153-234)
{
    if (mb.wide(25, 56) != 0) return 0;
    const uint32_t caps = mb(1, 24);                                           // capability bit k is bit (24 - k)
    auto has = [caps](int k) { return (caps >> (24 - k)) & 1u; };
    int score = has(7) ? 1 : -2;
    score -= 2 * int(has(10) + has(11) + has(12) + has(20) + has(21) + has(22)) + int(has(13) + has(14));
    const uint32_t es = caps >> 18;                                            // bits 1..6: the extended-squitter registers
    if ((es >> 1) == 0x1F) score += 5 + int(es & 1);
    else if (es == 0) score += 1;
    else if ((es >> 1) == 0x07) score += 3;
    else score -= 12;
    if (has(16) && has(24)) score += 2 + int(has(9));
    else if (!has(16) && !has(24) && !has(9)) score += 1;
    else score -= 6;
    return score;
}

__device__ __forceinline__ int score_ident20(const MeBits &mb, bool *usable)    // BDS 2,0 (comm_b.c:237-281)
{
    if (mb(1, 8) != 0x20) return 0;
    int score = 8;
    bool all_real = true;
    for (int k = 0; k < 8; ++k) {
        const uint32_t six = mb(9 + 6 * k, 14 + 6 * k);
        const bool letter = six >= 1 && six <= 26, digit = six >= 48 && six <= 57, space = six == 32;
        if (letter || digit || space) score += 6;
        else if (six == 0) all_real = false;                                   // '@' padding: still BDS 2,0, callsign not used
        else return 0;
    }
    *usable = all_real;
    return score;
}

// an optional field of BDS 4,0 / 5,0 / 6,0: status bit set and value plausible -> `gain`; status clear and field zero -> 1; else reject (-1)
__device__ __forceinline__ int field_score(bool status, bool nonzero, bool plausible, int gain)
{
    if (status && nonzero) return plausible ? gain : -1;
    if (!status && !nonzero) return 1;
    return -1;
}

struct Bds40 { uint32_t mcp_alt, fms_alt; float qnh; };

__device__ __forceinline__ int score_intent40(const MeBits &mb, Bds40 &d)       // BDS 4,0 (comm_b.c:302-465)
{
    const bool mcp_on = mb(1), fms_on = mb(14), baro_on = mb(27), mode_on = mb(48), source_on = mb(54);
    if (!mcp_on && !fms_on && !baro_on && !mode_on && !source_on) return 0;
    const uint32_t mcp_raw = mb(2, 13), fms_raw = mb(15, 26), baro_raw = mb(28, 39);
    d.mcp_alt = mcp_on ? mcp_raw * 16 : 0;
    d.fms_alt = fms_on ? fms_raw * 16 : 0;
    // float baro_setting = 800 + baro_raw * 0.1
    d.qnh = (baro_on && baro_raw) ? __double2float_rn(__dadd_rn(800.0, __dmul_rn(double(baro_raw), 0.1))) : 0.f;
    const int s_mcp = field_score(mcp_on, mcp_raw != 0, within(mcp_raw * 16, 1000, 50000), 13);
    const int s_fms = field_score(fms_on, fms_raw != 0, within(fms_raw * 16, 1000, 50000), 13);
    const int s_baro = field_score(baro_on, baro_raw != 0, d.qnh >= 900.f && d.qnh <= 1100.f, 13);
    const int s_mode = mode_on ? 4 : mb(49, 51) == 0 ? 1 : -1;
    const int s_source = source_on ? 3 : mb(55, 56) == 0 ? 1 : -1;
    if (s_mcp < 0 || s_fms < 0 || s_baro < 0 || s_mode < 0 || s_source < 0 || mb(40, 47) != 0 || mb(52, 53) != 0) return 0;
    int score = s_mcp + s_fms + s_baro + s_mode + s_source;
    auto off_grid = [](uint32_t alt) { const uint32_t r = alt % 500; return !(r < 16 || r > 484); };   // selected altitudes are multiples of 500 ft
    if (mcp_on && fms_on && d.mcp_alt != d.fms_alt) score -= 4;
    if (mcp_on && off_grid(d.mcp_alt)) score -= 4;
    if (fms_on && off_grid(d.fms_alt)) score -= 4;
    return score;
}

struct Bds50 { float roll, track, track_rate; uint32_t gs, tas; bool rate_on; };

__device__ __forceinline__ int score_track50(const MeBits &mb, Bds50 &d)        // BDS 5,0 (comm_b.c:468-623)
{
    if (!mb(1) || !mb(12) || !mb(24) || !mb(46)) return 0;                     // roll, track, ground speed and TAS must all be there
    // all of these are exact in float: k * 45/256 - 90, k * 90/512 + 180, k * 8/256 - 16
    d.roll = float(mb(3, 11)) * (45.f / 256.f) - (mb(2) ? 90.f : 0.f);
    d.track = float(mb(14, 23)) * (90.f / 512.f) + (mb(13) ? 180.f : 0.f);
    d.gs = mb(25, 34) * 2;
    d.tas = mb(47, 56) * 2;
    d.rate_on = mb(35);
    d.track_rate = d.rate_on ? float(mb(37, 45)) * (8.f / 256.f) - (mb(36) ? 16.f : 0.f) : 0.f;
    if (!(d.roll >= -40.f && d.roll < 40.f) || !within(d.gs, 50, 700) || !within(d.tas, 50, 700)) return 0;
    int score = 11 + 12 + 11 + 11;
    if (d.rate_on) {
        if (!(d.track_rate >= -10.f && d.track_rate <= 10.f)) return 0;
        score += 11;
        // the turn rate the bank angle implies at this airspeed against the reported one (comm_b.c:584-590); the reference's
        // ground speed / airspeed consistency test two lines earlier compares the status bits and never fires
        const double implied = __ddiv_rn(__dmul_rn(68625.0, tan(__ddiv_rn(__dmul_rn(double(d.roll), 3.14159265358979323846), 180.0))),
                                         __dmul_rn(double(d.tas * 20), 3.14159265358979323846));
        if (fabs(__dsub_rn(implied, double(d.track_rate))) > 2.0) score -= 6;
    } else {
        if (mb(36, 45) != 0) return 0;
        score += 1;
    }
    return score;
}

struct Bds60 { float heading, mach; uint32_t ias; int baro_rate, ins_rate; bool baro_on, ins_on; };

__device__ __forceinline__ int score_speed60(const MeBits &mb, Bds60 &d)        // BDS 6,0 (comm_b.c:626-775)
{
    d.baro_on = mb(35);
    d.ins_on = mb(46);
    if (!mb(1) || !mb(13) || !mb(24) || (!d.baro_on && !d.ins_on)) return 0;
    d.heading = float(mb(3, 12)) * (90.f / 512.f) + (mb(2) ? 180.f : 0.f);
    d.ias = mb(14, 23);
    const uint32_t mach_raw = mb(25, 34);
    d.mach = __double2float_rn(__ddiv_rn(__dmul_rn(double(mach_raw), 2.048), 512.0));      // float mach = mach_raw * 2.048 / 512
    if (!within(d.ias, 50, 700) || mach_raw == 0 || !(double(d.mach) >= 0.1 && double(d.mach) <= 0.9)) return 0;
    int score = 12 + 11 + 11;
    d.baro_rate = d.ins_rate = 0;
    if (d.baro_on) {
        d.baro_rate = int(mb(37, 45)) * 32 - (mb(36) ? 16384 : 0);
        if (d.baro_rate < -6000 || d.baro_rate > 6000) return 0;
        score += 11;
    } else {
        if (mb(37, 45) != 0) return 0;
        score += 1;
    }
    if (d.ins_on) {
        d.ins_rate = int(mb(48, 56)) * 32 - (mb(47) ? 16384 : 0);
        if (d.ins_rate < -6000 || d.ins_rate > 6000) return 0;
        score += 11;
    } else {
        if (mb(48, 56) != 0) return 0;
        score += 1;
    }
    if (d.baro_on && d.ins_on && abs(d.baro_rate - d.ins_rate) > 2000) score -= 12;
    return score;
}

__device__ __forceinline__ int score_mrar44(const MeBits &mb)                    // BDS 4,4 (comm_b.c:778-937)
{
    const uint32_t source = mb(1, 4), wind = mb(6, 14), asp = mb(36, 46);
    if (source == 0 || source >= 5 || !mb(5) || !mb(24)) return 0;             // wind and temperature are in every real report
    if ((!mb(35) && asp) || (!mb(47) && mb(48, 49)) || (!mb(50) && mb(51, 56))) return 0;
    const int quarter_deg = int(mb(26, 34)) - (mb(25) ? 512 : 0);              // static air temperature in units of 0.25 C
    int score;
    if (wind == 0) score = 2; else if (wind <= 250) score = 19; else return 0;
    if (quarter_deg == 0) score += 2; else if (quarter_deg >= -320 && quarter_deg <= 240) score += 11; else return 0;
    if (mb(35)) { if (!within(asp, 25, 1100)) return 0; score += 12; } else score += 1;
    score += (mb(47) ? 3 : 1) + (mb(50) ? 7 : 1);
    return source == 3 ? 1 : score;                                            // DME/DME look-alikes of a capability report: last resort
}

__device__ __forceinline__ int score_position05(const MeBits &mb, int df, uint32_t ac13)   // BDS 0,5 (comm_b.c:944-985)
{
    const uint32_t tc = mb(1, 5), ac12 = mb(9, 20);
    if (df != 20 || tc < 9 || tc > 18 || mb(21) || !ac12) return 0;
    if (ac13 != (((ac12 & 0x0FC0) << 1) | (ac12 & 0x003F))) return 0;          // must repeat the altitude of the reply around it
    return (mb(23, 39) == 0 || mb(40, 56) == 0) ? 0 : 100;
}

__device__ void commb_decode_one(const uint8_t *m, int df, int corrected, adsbcu_fields_commb &o)
{
    memset(&o, 0, sizeof(o));
    if (df != 20 && df != 21) return;
    MeBits mb;
    mb.v = 0;
    for (int i = 0; i < 7; ++i) mb.v = (mb.v << 8) | m[4 + i];
    // DR is bits 9-13 of the reply; the reference's UM test reads a field it has not filled in yet (mode_s.c:731-734 vs :757-760)
    const uint32_t dr = (uint32_t(m[1]) >> 3) & 0x1Fu;
    if (dr != 0 || corrected > 0) { o.commb_format = kCbNotDecoded; return; }
    const uint32_t ac13 = df == 20 ? (((uint32_t(m[2]) << 8) | m[3]) & 0x1FFFu) : 0u;
    const uint32_t bds = mb(1, 8);
    bool callsign_usable = false;
    Bds40 d40;
    Bds50 d50;
    Bds60 d60;
    int score[10];                                                             // the reference's order: a later equal score is a tie
    score[0] = ((bds == 0x00 || bds == 0x40 || bds == 0x50 || bds == 0x60) && mb.wide(9, 56) == 0) ? 56 : 0;      // empty response (comm_b.c:88-126)
    score[1] = (bds == 0x10 && mb(10, 14) == 0) ? 56 : 0;                      // BDS 1,0 (comm_b.c:129-150)
    score[2] = score_ident20(mb, &callsign_usable);
    score[3] = bds == 0x30 ? 56 : 0;                                           // BDS 3,0 (comm_b.c:284-299)
    score[4] = score_caps17(mb);
    score[5] = score_intent40(mb, d40);
    score[6] = score_track50(mb, d50);
    score[7] = score_speed60(mb, d60);
    score[8] = score_mrar44(mb);
    score[9] = score_position05(mb, df, ac13);
    int best = 0, winner = -1;
    bool tie = false;
#pragma unroll
    for (int k = 0; k < 10; ++k) {
        if (score[k] > best) { best = score[k]; winner = k; tie = false; }
        else if (score[k] == best) tie = true;
    }
    if (winner < 0) { o.commb_format = kCbUnknown; return; }
    if (tie) { o.commb_format = kCbAmbiguous; return; }
    uint32_t valid = 0;
    switch (winner) {
    case 0: o.commb_format = kCbEmpty; break;
    case 1: o.commb_format = kCbDatalinkCaps; break;
    case 2:
        o.commb_format = kCbIdent;
        if (callsign_usable) {
            valid |= ADSBCU_MB_CALLSIGN;
            for (int k = 0; k < 8; ++k) { const uint32_t six = mb(9 + 6 * k, 14 + 6 * k); o.callsign[k] = char(six < 32 ? 0x40 + six : six); }
        }
        break;
    case 3: o.commb_format = kCbAcasRa; break;
    case 4: o.commb_format = kCbGicbCaps; break;
    case 5: {
        o.commb_format = kCbVerticalIntent;
        if (mb(1)) { valid |= ADSBCU_MB_NAV_MCP_ALTITUDE; o.nav_mcp_altitude = int(d40.mcp_alt); }
        if (mb(14)) { valid |= ADSBCU_MB_NAV_FMS_ALTITUDE; o.nav_fms_altitude = int(d40.fms_alt); }
        if (mb(27)) { valid |= ADSBCU_MB_NAV_QNH; o.nav_qnh = d40.qnh; }
        if (mb(48)) {
            const uint32_t mode = mb(49, 51);
            valid |= ADSBCU_MB_NAV_MODES;
            o.nav_modes = uint8_t(((mode & 4) ? kNavVnav : 0) | ((mode & 2) ? kNavAltHold : 0) | ((mode & 1) ? kNavApproach : 0));
        }
        o.nav_altitude_source = mb(54) ? uint8_t(1 + mb(55, 56)) : 0;          // NAV_ALT_UNKNOWN / AIRCRAFT / MCP / FMS are 1..4
        break;
    }
    case 6:
        o.commb_format = kCbTrackTurn;
        valid |= ADSBCU_MB_ROLL | ADSBCU_MB_HEADING | ADSBCU_MB_GS | ADSBCU_MB_TAS;
        o.roll = d50.roll;
        o.heading = d50.track;
        o.heading_type = kHdgGroundTrack;
        o.gs = float(d50.gs);
        o.tas = uint16_t(d50.tas);
        if (d50.rate_on) { valid |= ADSBCU_MB_TRACK_RATE; o.track_rate = d50.track_rate; }
        break;
    case 7:
        o.commb_format = kCbHeadingSpeed;
        valid |= ADSBCU_MB_HEADING | ADSBCU_MB_IAS | ADSBCU_MB_MACH;
        o.heading = d60.heading;
        o.heading_type = kHdgMagnetic;
        o.ias = uint16_t(d60.ias);
        o.mach = d60.mach;
        if (d60.baro_on) { valid |= ADSBCU_MB_BARO_RATE; o.baro_rate = d60.baro_rate; }
        if (d60.ins_on) { valid |= ADSBCU_MB_GEOM_RATE; o.geom_rate = d60.ins_rate; }       // inertial rate is filed as the geometric one
        break;
    case 8:
        o.commb_format = kCbMrar;
        valid |= ADSBCU_MB_MRAR_SOURCE | ADSBCU_MB_WIND | ADSBCU_MB_TEMPERATURE;
        o.mrar_source = uint8_t(mb(1, 4));
        o.wind_speed = float(mb(6, 14));
        o.wind_dir = float(mb(15, 23)) * (180.f / 256.f);
        o.temperature = float(int(mb(26, 34)) - (mb(25) ? 512 : 0)) * 0.25f;
        if (mb(35)) { valid |= ADSBCU_MB_PRESSURE; o.pressure = float(mb(36, 46)); }
        if (mb(47)) { valid |= ADSBCU_MB_TURBULENCE; o.turbulence = uint8_t(mb(48, 49)); }
        if (mb(50)) { valid |= ADSBCU_MB_HUMIDITY; o.humidity = float(mb(51, 56)) * (100.f / 64.f); }
        break;
    default: o.commb_format = kCbAirbornePosition; break;                      // recognised only, never used as a position
    }
    o.valid = valid;
}

__global__ void __launch_bounds__(kEsThreads) fields_commb_kernel(const uint32_t *msgs, uint32_t n, uint32_t *out)
{
    constexpr uint32_t kInWords = sizeof(adsbcu_msg) / 4, kOutWords = sizeof(adsbcu_fields_commb) / 4;
    __shared__ uint32_t s_in[kEsThreads * kInWords];
    __shared__ uint32_t s_out[kEsThreads * kOutWords];
    const uint32_t first = blockIdx.x * kEsThreads;
    const uint32_t count = min(uint32_t(kEsThreads), n - first);
    for (uint32_t w = threadIdx.x; w < count * kInWords; w += kEsThreads) s_in[w] = msgs[size_t(first) * kInWords + w];
    __syncthreads();
    if (threadIdx.x < count) {
        const uint8_t *rec = reinterpret_cast<const uint8_t *>(s_in + threadIdx.x * kInWords);
        adsbcu_fields_commb o;
        commb_decode_one(rec + 8, rec[8 + 14 + 14 + 2], rec[8 + 14 + 14 + 1], o);
        const uint32_t *src = reinterpret_cast<const uint32_t *>(&o);
        for (uint32_t w = 0; w < kOutWords; ++w) s_out[threadIdx.x * kOutWords + w] = src[w];
    }
    __syncthreads();
    for (uint32_t w = threadIdx.x; w < count * kOutWords; w += kEsThreads) out[size_t(first) * kOutWords + w] = s_out[w];
}

}  // namespace

int fields_commb_decode(Fields *&f, cudaStream_t s, const adsbcu_msg *msgs, size_t n, adsbcu_fields_commb *out, std::string &err)
{
    static_assert(sizeof(adsbcu_fields_commb) == 84 && sizeof(adsbcu_msg) == 56, "record layout");
    if (!f) f = new Fields();
    if (!n) return ADSBCU_OK;
    cudaError_t e;
    if (!grow(f->d_in, f->cap_in, n * sizeof(adsbcu_msg)) || !grow(f->d_out, f->cap_out, n * sizeof(adsbcu_fields_commb)))
        return fail(err, "cudaMalloc(fields)", cudaGetLastError());
    if ((e = cudaMemcpyAsync(f->d_in, msgs, n * sizeof(adsbcu_msg), cudaMemcpyHostToDevice, s)) != cudaSuccess) return fail(err, "H2D(commb)", e);
    ADSBCU_LAUNCH(fields_commb_kernel, unsigned((n + kEsThreads - 1) / kEsThreads), kEsThreads, 0, s, static_cast<const uint32_t *>(f->d_in),
                  uint32_t(n), static_cast<uint32_t *>(f->d_out));
    if ((e = cudaGetLastError()) != cudaSuccess) return fail(err, "fields_commb_kernel", e);
    if ((e = cudaMemcpyAsync(out, f->d_out, n * sizeof(adsbcu_fields_commb), cudaMemcpyDeviceToHost, s)) != cudaSuccess) return fail(err, "D2H(commb)", e);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(err, "commb sync", e);
    return ADSBCU_OK;
}

// ---- Path A: all of decodeModesMessage for records that are already on the device ---------------------------------------
// One launch fills up to three output arrays (header fields, extended squitter, Comm-B) from device-resident adsbcu_msg
// records: the input tile is staged once, every output goes through the same shared staging tile so that the global
// writes are whole words in order.  No allocation, no copy, no wait: the caller's stream orders it.
namespace {

// tile <-> global copies: 16 bytes per thread and instruction when the global side is 16-byte aligned (every full tile of a
// cudaMalloc'ed array is: 256 records of 56 / 32 / 120 / 84 bytes), words otherwise
__device__ __forceinline__ void copy_words(uint32_t *dst, const uint32_t *src, uint32_t nwords)
{
    if (((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15u) == 0) {
        const uint32_t nvec = nwords / 4;
        for (uint32_t v = threadIdx.x; v < nvec; v += kEsThreads) reinterpret_cast<uint4 *>(dst)[v] = reinterpret_cast<const uint4 *>(src)[v];
        for (uint32_t w = nvec * 4 + threadIdx.x; w < nwords; w += kEsThreads) dst[w] = src[w];
    } else {
        for (uint32_t w = threadIdx.x; w < nwords; w += kEsThreads) dst[w] = src[w];
    }
}

// decode into the thread's slot of the shared tile (the record never lives in registers), then write the tile out in order
template <class Rec, class Decode>
__device__ __forceinline__ void decode_tile(Decode decode, bool live, uint32_t *s_out, uint32_t *out, uint32_t first, uint32_t count)
{
    constexpr uint32_t kWords = sizeof(Rec) / 4;
    __syncthreads();                                                           // the tile may still be read by the previous output's copy
    if (live) decode(*reinterpret_cast<Rec *>(s_out + threadIdx.x * kWords));
    __syncthreads();
    copy_words(out + size_t(first) * kWords, s_out, count * kWords);
}

__global__ void __launch_bounds__(kEsThreads, 4) fields_all_kernel(const uint32_t *msgs, uint32_t n, uint32_t *out_a, uint32_t *out_es, uint32_t *out_mb)
{
    constexpr uint32_t kInWords = sizeof(adsbcu_msg) / 4;
    __shared__ __align__(16) uint32_t s_in[kEsThreads * kInWords];
    __shared__ __align__(16) uint32_t s_out[kEsThreads * sizeof(adsbcu_fields_es) / 4];      // the largest of the three records
    for (uint32_t first = blockIdx.x * kEsThreads; first < n; first += gridDim.x * kEsThreads) {
        const uint32_t count = min(uint32_t(kEsThreads), n - first);
        __syncthreads();
        copy_words(s_in, msgs + size_t(first) * kInWords, count * kInWords);
        __syncthreads();
        const bool live = threadIdx.x < count;
        const uint8_t *rec = reinterpret_cast<const uint8_t *>(s_in + (live ? threadIdx.x : 0) * kInWords);
        const int df = rec[8 + 14 + 14 + 2], corrected = rec[8 + 14 + 14 + 1];
        if (out_a) decode_tile<adsbcu_fields_a>([&](adsbcu_fields_a &a) { fields_a_one(rec + 8, df, a); }, live, s_out, out_a, first, count);
        if (out_es) decode_tile<adsbcu_fields_es>([&](adsbcu_fields_es &e) { es_decode_one(rec + 8, df, corrected, e); }, live, s_out, out_es, first, count);
        if (out_mb) decode_tile<adsbcu_fields_commb>([&](adsbcu_fields_commb &b) { commb_decode_one(rec + 8, df, corrected, b); }, live, s_out, out_mb, first, count);
    }
}

}  // namespace

int fields_all_device(int sm_count, cudaStream_t s, const adsbcu_msg *msgs_dev, size_t n, adsbcu_fields_a *a_dev, adsbcu_fields_es *es_dev,
                      adsbcu_fields_commb *mb_dev, std::string &err)
{
    if (!n) return ADSBCU_OK;
    const unsigned tiles = unsigned((n + kEsThreads - 1) / kEsThreads);
    const unsigned grid = std::min(tiles, unsigned(sm_count) * 4u);            // 4 resident CTAs per SM (44 KiB of shared memory each)
    ADSBCU_LAUNCH(fields_all_kernel, grid, kEsThreads, 0, s, reinterpret_cast<const uint32_t *>(msgs_dev), uint32_t(n),
                  reinterpret_cast<uint32_t *>(a_dev), reinterpret_cast<uint32_t *>(es_dev), reinterpret_cast<uint32_t *>(mb_dev));
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(err, "fields_all_kernel", e);
    return ADSBCU_OK;
}

void fields_free(Fields *f)
{
    if (!f) return;
    cudaFree(f->d_in);
    cudaFree(f->d_out);
    cudaFree(f->d_heading);
    delete f;
}

}  // namespace adsbcu
