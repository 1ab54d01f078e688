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

__global__ void __launch_bounds__(256) fields_a_kernel(const adsbcu_msg *msgs, uint32_t n, adsbcu_fields_a *out)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    uint8_t m[14];
    const uint8_t *src = reinterpret_cast<const uint8_t *>(&msgs[k]) + 8;        // adsbcu_msg::msg (packed record: byte loads)
    for (int i = 0; i < 14; ++i) m[i] = src[i];
    const int df = reinterpret_cast<const uint8_t *>(&msgs[k])[8 + 14 + 14 + 2];   // adsbcu_msg::msgtype
    adsbcu_fields_a f;
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

void fields_free(Fields *f)
{
    if (!f) return;
    cudaFree(f->d_in);
    cudaFree(f->d_out);
    cudaFree(f->d_heading);
    delete f;
}

}  // namespace adsbcu
