/*
 * TEST INFRASTRUCTURE ONLY -- never linked into or called from the product path.
 *
 * CPU restatement of the CPR position decoders of the flightaware dump1090 tree the reference vendors
 * (dump1090/cpr.c): global airborne (:162-214), global surface (:216-311), receiver/aircraft-relative
 * (:325-369), with the NL zone table of 1090-WP-9-14 (:77-138).  SURVEY 8f rank 2, Path A side.
 *
 * Pinned (tests/test_cpr.py) against
 *   - the golden vectors of the reference's own dump1090/cprtests.c (tests/golden/cpr_vectors.npz), and
 *   - oracle/_ref/libcpr_ref.so = the reference's cpr.c compiled where it lies, on those vectors and on random
 *     inputs, bit for bit (the outputs are IEEE doubles; no -ffast-math, no FMA contraction: x86-64 baseline).
 *
 * One record in, one record out, mirroring what the three reference functions take and return:
 *   mode 0  decodeCPRairborne(even_lat, even_lon, odd_lat, odd_lon, fflag)
 *   mode 1  decodeCPRsurface(reflat, reflon, even_lat, even_lon, odd_lat, odd_lon, fflag)
 *   mode 2  decodeCPRrelative(reflat, reflon, cprlat = even_lat, cprlon = even_lon, fflag, surface = 0)
 *   mode 3  decodeCPRrelative(..., surface = 1)
 * result is the reference's return value (0 ok, -1, -2); lat/lon are untouched (0) unless result == 0.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>

typedef struct {            /* == adsbcu_cpr_in (include/adsbcu.h) */
    double  reflat, reflon;
    int32_t even_lat, even_lon, odd_lat, odd_lon;
    int32_t fflag, mode;
} cpr_in;
typedef struct {            /* == adsbcu_cpr_out */
    double  lat, lon;
    int32_t result, reserved;
} cpr_out;

static const double nl_edge[58] = {     /* cpr.c:79-136: NL = 59 - (number of edges <= |lat|) */
    10.47047130, 14.82817437, 18.18626357, 21.02939493, 23.54504487, 25.82924707, 27.93898710, 29.91135686, 31.77209708,
    33.53993436, 35.22899598, 36.85025108, 38.41241892, 39.92256684, 41.38651832, 42.80914012, 44.19454951, 45.54626723,
    46.86733252, 48.16039128, 49.42776439, 50.67150166, 51.89342469, 53.09516153, 54.27817472, 55.44378444, 56.59318756,
    57.72747354, 58.84763776, 59.95459277, 61.04917774, 62.13216659, 63.20427479, 64.26616523, 65.31845310, 66.36171008,
    67.39646774, 68.42322022, 69.44242631, 70.45451075, 71.45986473, 72.45884545, 73.45177442, 74.43893416, 75.42056257,
    76.39684391, 77.36789461, 78.33374083, 79.29428225, 80.24923213, 81.19801349, 82.13956981, 83.07199445, 83.99173563,
    84.89166191, 85.75541621, 86.53536998, 87.00000000};

static int nl(double lat)
{
    int k = 0;
    lat = fabs(lat);
    while (k < 58 && !(lat < nl_edge[k])) ++k;
    return 59 - k;
}
static int zones(double lat, int odd) { const int n = nl(lat) - (odd ? 1 : 0); return n < 1 ? 1 : n; }   /* cpr.c:142-146 */
static int pmod(int a, int b) { const int r = a % b; return r < 0 ? r + b : r; }
static double pfmod(double a, double b) { const double r = fmod(a, b); return r < 0 ? r + b : r; }

/* surface quadrant choice for one latitude candidate (cpr.c:262-278) */
static double surface_quadrant(double rlat, double reflat)
{
    if (rlat == 0) return reflat < -45 ? -90 : reflat > 45 ? 90 : rlat;
    return (rlat - reflat) > 45 ? rlat - 90 : rlat;
}

/* the two global decodes share everything but the span (360 or 90 degrees), the latitude fix-up and the
 * longitude quadrant step */
static int global_decode(const cpr_in *in, int surface, cpr_out *out)
{
    const double span = surface ? 90.0 : 360.0;
    const double lat0 = in->even_lat, lat1 = in->odd_lat, lon0 = in->even_lon, lon1 = in->odd_lon;
    const int j = (int) floor(((59 * lat0 - 60 * lat1) / 131072) + 0.5);
    double rlat0 = (span / 60.0) * (pmod(j, 60) + lat0 / 131072);
    double rlat1 = (span / 59.0) * (pmod(j, 59) + lat1 / 131072);
    if (surface) {
        rlat0 = surface_quadrant(rlat0, in->reflat);
        rlat1 = surface_quadrant(rlat1, in->reflat);
    } else {
        if (rlat0 >= 270) rlat0 -= 360;
        if (rlat1 >= 270) rlat1 -= 360;
    }
    if (rlat0 < -90 || rlat0 > 90 || rlat1 < -90 || rlat1 > 90) return -2;
    if (nl(rlat0) != nl(rlat1)) return -1;

    const int odd = in->fflag != 0;
    const double rlat = odd ? rlat1 : rlat0;
    const int n = nl(rlat), ni = zones(rlat, odd);
    const int m = (int) floor((((lon0 * (n - 1)) - (lon1 * n)) / 131072) + 0.5);
    double rlon = (span / ni) * (pmod(m, ni) + (odd ? lon1 : lon0) / 131072);
    if (surface) rlon += floor((in->reflon - rlon + 45) / 90) * 90;     /* towards the reference, cpr.c:303 */
    rlon -= floor((rlon + 180) / 360) * 360;
    out->lat = rlat;
    out->lon = rlon;
    return 0;
}

static int relative_decode(const cpr_in *in, int surface, cpr_out *out)     /* cpr.c:325-369 */
{
    const int odd = in->fflag != 0;
    const double flat = in->even_lat / 131072.0, flon = in->even_lon / 131072.0;
    const double dlat = (surface ? 90.0 : 360.0) / (odd ? 59.0 : 60.0);
    const int j = (int) (floor(in->reflat / dlat) + floor(0.5 + pfmod(in->reflat, dlat) / dlat - flat));
    double rlat = dlat * (j + flat);
    if (rlat >= 270) rlat -= 360;
    if (rlat < -90 || rlat > 90) return -1;
    if (fabs(rlat - in->reflat) > (dlat / 2)) return -1;
    const double dlon = (surface ? 90.0 : 360.0) / zones(rlat, odd);
    const int m = (int) (floor(in->reflon / dlon) + floor(0.5 + pfmod(in->reflon, dlon) / dlon - flon));
    double rlon = dlon * (m + flon);
    if (rlon > 180) rlon -= 360;
    if (fabs(rlon - in->reflon) > (dlon / 2)) return -1;
    out->lat = rlat;
    out->lon = rlon;
    return 0;
}

void oraclecpr_decode(const cpr_in *in, size_t n, cpr_out *out)
{
    for (size_t k = 0; k < n; ++k) {
        out[k].lat = out[k].lon = 0;
        out[k].reserved = 0;
        switch (in[k].mode) {
        case 0: out[k].result = global_decode(&in[k], 0, &out[k]); break;
        case 1: out[k].result = global_decode(&in[k], 1, &out[k]); break;
        case 2: out[k].result = relative_decode(&in[k], 0, &out[k]); break;
        default: out[k].result = relative_decode(&in[k], 1, &out[k]); break;
        }
    }
}
int oraclecpr_nl(double lat) { return nl(lat); }
