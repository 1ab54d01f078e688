/*
 * TEST INFRASTRUCTURE ONLY -- not part of the shipped product path.
 *
 * Harness linked into oracle/_ref/dump1090_ref, the UNMODIFIED flightaware
 * dump1090 compiled from the sources where they lie under
 * /root/reference/dump1090 (see oracle/Makefile).  It adds exactly three things:
 *
 *  1. the four interactive* entry points declared in dump1090.h (the real
 *     interactive.c needs <curses.h>, absent in this image); all no-ops, never
 *     reached in --ifile batch mode.
 *  2. a link-time wrap of useModesMessage() (mode_s.c), the sink that
 *     demodulate2400() hands every accepted message to (demod_2400.c:422).
 *     When REF1090_OUT is set the wrap appends one packed binary record per
 *     message (the parity tuple: bytes, verbatim, crc, corrected bits,
 *     12 MHz timestamp, score, addr, RSSI) and then calls the real sink, so
 *     the reference's own behaviour (tracking, stats, printing) is unchanged.
 *  3. a link-time wrap of main(): with REF1090_DECODE_IN / REF1090_DECODE_OUT set the binary does not demodulate
 *     anything; it reads 14-byte frames, hands each to the reference's decodeModesMessage() (mode_s.c:520) exactly
 *     as demodulate2400 does (zeroed struct, score from scoreModesMessage) and writes what the extended-squitter
 *     and Comm-B decoders left in struct modesMessage, in the layouts of adsbcu_fields_es and adsbcu_fields_commb, for
 *     tests/make_golden_es.py and tests/make_golden_mb.py.
 */
#include "dump1090.h"
#include "adsbcu.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

void interactiveInit(void) {}
void interactiveShowData(void) {}
void interactiveCleanup(void) {}
void interactiveNoConnection(void) {}

void __real_useModesMessage(struct modesMessage *mm);

#pragma pack(push, 1)
struct ref_record {
    uint64_t timestamp;      /* mm->timestampMsg (12 MHz) */
    uint8_t  msg[14];        /* corrected bytes */
    uint8_t  verbatim[14];   /* as sliced */
    uint8_t  msgbits;        /* 56 / 112 */
    uint8_t  correctedbits;
    uint8_t  msgtype;        /* DF */
    uint8_t  score;
    uint32_t crc;
    uint32_t addr;
    double   signal_level;
};
#pragma pack(pop)

static FILE *out_fp;
static int   out_checked;

static void close_out(void)
{
    if (out_fp) fclose(out_fp);
    out_fp = NULL;
}

/* the surveillance fields decodeModesMessage filled in (mode_s.c:626-779), in the layout of adsbcu_fields_a */
struct hdr_fields { uint16_t AC, ID; uint8_t CA, CC, CF, DR, FS, KE, ND, RI, SL, UM, VS, airground, alert_valid, alert, spi_valid, spi,
                    altitude_baro_valid, altitude_baro_unit, squawk_valid, es_pending; uint16_t squawk; uint16_t pad; int32_t altitude_baro; };
static void hdr_record(const struct modesMessage *mm, struct hdr_fields *out)
{
    struct hdr_fields f;
    memset(&f, 0, sizeof(f));
    f.AC = (uint16_t) mm->AC; f.ID = (uint16_t) mm->ID; f.CA = (uint8_t) mm->CA; f.CC = (uint8_t) mm->CC; f.CF = (uint8_t) mm->CF;
    f.DR = (uint8_t) mm->DR; f.FS = (uint8_t) mm->FS; f.KE = (uint8_t) mm->KE; f.ND = (uint8_t) mm->ND; f.RI = (uint8_t) mm->RI;
    f.SL = (uint8_t) mm->SL; f.UM = (uint8_t) mm->UM; f.VS = (uint8_t) mm->VS; f.airground = (uint8_t) mm->airground;
    f.alert_valid = mm->alert_valid; f.alert = mm->alert; f.spi_valid = mm->spi_valid; f.spi = mm->spi;
    f.altitude_baro_valid = mm->altitude_baro_valid; f.altitude_baro_unit = (uint8_t) mm->altitude_baro_unit;
    f.squawk_valid = mm->squawk_valid; f.es_pending = (mm->msgtype == 17 || mm->msgtype == 18);
    f.squawk = (uint16_t) mm->squawk; f.altitude_baro = mm->altitude_baro;
    *out = f;
}

void __wrap_useModesMessage(struct modesMessage *mm)
{
    if (!out_checked) {
        const char *p = getenv("REF1090_OUT");
        out_checked = 1;
        if (p && *p) {
            out_fp = fopen(p, "wb");
            if (out_fp) {
                setvbuf(out_fp, NULL, _IOFBF, 1 << 20);
                atexit(close_out);
            }
        }
    }
    if (out_fp) {
        struct ref_record r;
        memset(&r, 0, sizeof(r));
        r.timestamp = mm->timestampMsg;
        memcpy(r.msg, mm->msg, 14);
        memcpy(r.verbatim, mm->verbatim, 14);
        r.msgbits = (uint8_t) mm->msgbits;
        r.correctedbits = (uint8_t) mm->correctedbits;
        r.msgtype = (uint8_t) mm->msgtype;
        r.score = (uint8_t) mm->score;
        r.crc = mm->crc;
        r.addr = mm->addr;
        r.signal_level = mm->signalLevel;
        fwrite(&r, sizeof(r), 1, out_fp);
    }
    {
        /* second tap (REF1090_FIELDS=path): the surveillance fields decodeModesMessage filled in (mode_s.c:626-779), in the
         * layout of adsbcu_fields_a, for tests/make_golden_fields.py */
        static FILE *ffp;
        static int fchecked;
        if (!fchecked) {
            const char *p = getenv("REF1090_FIELDS");
            fchecked = 1;
            if (p && *p) ffp = fopen(p, "wb");
        }
        if (ffp) {
            struct hdr_fields f;
            hdr_record(mm, &f);
            fwrite(&f, sizeof(f), 1, ffp);
            fflush(ffp);
        }
    }
    __real_useModesMessage(mm);
}

/* ---- third tap: decodeModesMessage on frames from a file ------------------------------------------------------- */
int __real_main(int argc, char **argv);

static void es_record(const struct modesMessage *mm, adsbcu_fields_es *e)
{
    unsigned v = 0, o = 0;
    memset(e, 0, sizeof(*e));
    if (mm->msgtype != 17 && mm->msgtype != 18) return;
    e->es = 1;
    e->addr = mm->addr;
    e->metype = (uint8_t) mm->metype; e->mesub = (uint8_t) mm->mesub; e->addrtype = (uint8_t) mm->addrtype;
    e->source = (uint8_t) mm->source; e->reliable = (uint8_t) mm->reliable; e->airground = (uint8_t) mm->airground;
    e->CA_CF = (uint8_t) (mm->msgtype == 17 ? mm->CA : mm->CF);
    e->cpr_type = (uint8_t) mm->cpr_type; e->cpr_odd = mm->cpr_odd; e->heading_type = (uint8_t) mm->heading_type;
    e->nav_heading_type = (uint8_t) mm->nav.heading_type;
    e->alert = mm->alert; e->spi = mm->spi; e->emergency = (uint8_t) mm->emergency; e->category = (uint8_t) mm->category;
    e->altitude_baro_unit = (uint8_t) mm->altitude_baro_unit; e->altitude_geom_unit = (uint8_t) mm->altitude_geom_unit;
    e->nav_altitude_source = (uint8_t) mm->nav.altitude_source; e->nav_modes = (uint8_t) mm->nav.modes;
    memcpy(e->callsign, mm->callsign, 8);
    e->altitude_baro = mm->altitude_baro; e->altitude_geom = mm->altitude_geom; e->geom_delta = mm->geom_delta;
    e->baro_rate = mm->baro_rate; e->geom_rate = mm->geom_rate;
    e->cpr_lat = mm->cpr_lat; e->cpr_lon = mm->cpr_lon;
    e->ias = (uint16_t) mm->ias; e->tas = (uint16_t) mm->tas; e->squawk = (uint16_t) mm->squawk;
    e->heading = mm->heading; e->gs_v0 = mm->gs.v0; e->gs_v2 = mm->gs.v2; e->gs_selected = mm->gs.selected;
    e->nav_heading = mm->nav.heading; e->nav_qnh = mm->nav.qnh;
    e->nav_fms_altitude = mm->nav.fms_altitude; e->nav_mcp_altitude = mm->nav.mcp_altitude;
    e->nic_a = mm->accuracy.nic_a; e->nic_b = mm->accuracy.nic_b; e->nic_c = mm->accuracy.nic_c; e->nic_baro = mm->accuracy.nic_baro;
    e->nac_p = (uint8_t) mm->accuracy.nac_p; e->nac_v = (uint8_t) mm->accuracy.nac_v; e->sil = (uint8_t) mm->accuracy.sil;
    e->sil_type = (uint8_t) mm->accuracy.sil_type; e->gva = (uint8_t) mm->accuracy.gva; e->sda = (uint8_t) mm->accuracy.sda;
    e->ops_version = (uint8_t) mm->opstatus.version; e->ops_tah = (uint8_t) mm->opstatus.tah; e->ops_hrd = (uint8_t) mm->opstatus.hrd;
    e->ops_cc_tc = (uint8_t) mm->opstatus.cc_tc; e->ops_cc_lw = (uint8_t) mm->opstatus.cc_lw;
    e->ops_cc_antenna_offset = (uint8_t) mm->opstatus.cc_antenna_offset;
    if (mm->altitude_baro_valid) v |= ADSBCU_ES_ALTITUDE_BARO;
    if (mm->altitude_geom_valid) v |= ADSBCU_ES_ALTITUDE_GEOM;
    if (mm->heading_valid) v |= ADSBCU_ES_HEADING;
    if (mm->gs_valid) v |= ADSBCU_ES_GS;
    if (mm->ias_valid) v |= ADSBCU_ES_IAS;
    if (mm->tas_valid) v |= ADSBCU_ES_TAS;
    if (mm->baro_rate_valid) v |= ADSBCU_ES_BARO_RATE;
    if (mm->geom_rate_valid) v |= ADSBCU_ES_GEOM_RATE;
    if (mm->squawk_valid) v |= ADSBCU_ES_SQUAWK;
    if (mm->callsign_valid) v |= ADSBCU_ES_CALLSIGN;
    if (mm->cpr_valid) v |= ADSBCU_ES_CPR;
    if (mm->category_valid) v |= ADSBCU_ES_CATEGORY;
    if (mm->geom_delta_valid) v |= ADSBCU_ES_GEOM_DELTA;
    if (mm->spi_valid) v |= ADSBCU_ES_SPI;
    if (mm->alert_valid) v |= ADSBCU_ES_ALERT;
    if (mm->emergency_valid) v |= ADSBCU_ES_EMERGENCY;
    if (mm->accuracy.nic_a_valid) v |= ADSBCU_ES_NIC_A;
    if (mm->accuracy.nic_b_valid) v |= ADSBCU_ES_NIC_B;
    if (mm->accuracy.nic_c_valid) v |= ADSBCU_ES_NIC_C;
    if (mm->accuracy.nic_baro_valid) v |= ADSBCU_ES_NIC_BARO;
    if (mm->accuracy.nac_p_valid) v |= ADSBCU_ES_NAC_P;
    if (mm->accuracy.nac_v_valid) v |= ADSBCU_ES_NAC_V;
    if (mm->accuracy.gva_valid) v |= ADSBCU_ES_GVA;
    if (mm->accuracy.sda_valid) v |= ADSBCU_ES_SDA;
    if (mm->nav.heading_valid) v |= ADSBCU_ES_NAV_HEADING;
    if (mm->nav.fms_altitude_valid) v |= ADSBCU_ES_NAV_FMS_ALTITUDE;
    if (mm->nav.mcp_altitude_valid) v |= ADSBCU_ES_NAV_MCP_ALTITUDE;
    if (mm->nav.qnh_valid) v |= ADSBCU_ES_NAV_QNH;
    if (mm->nav.modes_valid) v |= ADSBCU_ES_NAV_MODES;
    if (mm->opstatus.valid) v |= ADSBCU_ES_OPSTATUS;
    if (mm->opstatus.cc_lw_valid) v |= ADSBCU_ES_CC_LW;
    e->valid = v;
    if (mm->opstatus.om_acas_ra) o |= ADSBCU_OPS_OM_ACAS_RA;
    if (mm->opstatus.om_ident) o |= ADSBCU_OPS_OM_IDENT;
    if (mm->opstatus.om_atc) o |= ADSBCU_OPS_OM_ATC;
    if (mm->opstatus.om_saf) o |= ADSBCU_OPS_OM_SAF;
    if (mm->opstatus.cc_acas) o |= ADSBCU_OPS_CC_ACAS;
    if (mm->opstatus.cc_cdti) o |= ADSBCU_OPS_CC_CDTI;
    if (mm->opstatus.cc_1090_in) o |= ADSBCU_OPS_CC_1090_IN;
    if (mm->opstatus.cc_arv) o |= ADSBCU_OPS_CC_ARV;
    if (mm->opstatus.cc_ts) o |= ADSBCU_OPS_CC_TS;
    if (mm->opstatus.cc_uat_in) o |= ADSBCU_OPS_CC_UAT_IN;
    if (mm->opstatus.cc_poa) o |= ADSBCU_OPS_CC_POA;
    if (mm->opstatus.cc_b2_low) o |= ADSBCU_OPS_CC_B2_LOW;
    e->ops_flags = (uint16_t) o;
}

static void mb_record(const struct modesMessage *mm, adsbcu_fields_commb *b)
{
    unsigned v = 0;
    memset(b, 0, sizeof(*b));
    if (mm->msgtype != 20 && mm->msgtype != 21) return;
    b->commb_format = (uint8_t) mm->commb_format; b->heading_type = (uint8_t) mm->heading_type;
    b->nav_modes = (uint8_t) mm->nav.modes; b->nav_altitude_source = (uint8_t) mm->nav.altitude_source;
    b->mrar_source = (uint8_t) mm->mrar_source; b->turbulence = (uint8_t) mm->turbulence;
    memcpy(b->callsign, mm->callsign, 8);
    b->roll = mm->roll; b->heading = mm->heading; b->track_rate = mm->track_rate; b->gs = mm->gs.selected; b->mach = (float) mm->mach;
    b->nav_qnh = mm->nav.qnh; b->wind_speed = mm->wind_speed; b->wind_dir = mm->wind_dir; b->temperature = mm->temperature;
    b->pressure = mm->pressure; b->humidity = mm->humidity;
    b->baro_rate = mm->baro_rate; b->geom_rate = mm->geom_rate; b->nav_mcp_altitude = mm->nav.mcp_altitude; b->nav_fms_altitude = mm->nav.fms_altitude;
    b->ias = (uint16_t) mm->ias; b->tas = (uint16_t) mm->tas;
    if (mm->callsign_valid) v |= ADSBCU_MB_CALLSIGN;
    if (mm->roll_valid) v |= ADSBCU_MB_ROLL;
    if (mm->heading_valid) v |= ADSBCU_MB_HEADING;
    if (mm->gs_valid) v |= ADSBCU_MB_GS;
    if (mm->track_rate_valid) v |= ADSBCU_MB_TRACK_RATE;
    if (mm->tas_valid) v |= ADSBCU_MB_TAS;
    if (mm->ias_valid) v |= ADSBCU_MB_IAS;
    if (mm->mach_valid) v |= ADSBCU_MB_MACH;
    if (mm->baro_rate_valid) v |= ADSBCU_MB_BARO_RATE;
    if (mm->geom_rate_valid) v |= ADSBCU_MB_GEOM_RATE;
    if (mm->nav.mcp_altitude_valid) v |= ADSBCU_MB_NAV_MCP_ALTITUDE;
    if (mm->nav.fms_altitude_valid) v |= ADSBCU_MB_NAV_FMS_ALTITUDE;
    if (mm->nav.qnh_valid) v |= ADSBCU_MB_NAV_QNH;
    if (mm->nav.modes_valid) v |= ADSBCU_MB_NAV_MODES;
    if (mm->mrar_source_valid) v |= ADSBCU_MB_MRAR_SOURCE;
    if (mm->wind_valid) v |= ADSBCU_MB_WIND;
    if (mm->temperature_valid) v |= ADSBCU_MB_TEMPERATURE;
    if (mm->pressure_valid) v |= ADSBCU_MB_PRESSURE;
    if (mm->turbulence_valid) v |= ADSBCU_MB_TURBULENCE;
    if (mm->humidity_valid) v |= ADSBCU_MB_HUMIDITY;
    b->valid = v;
}

int __wrap_main(int argc, char **argv)
{
    const char *in = getenv("REF1090_DECODE_IN"), *out = getenv("REF1090_DECODE_OUT");
    if (!in || !out || !*in || !*out) return __real_main(argc, argv);
    FILE *fi = fopen(in, "rb"), *fo = fopen(out, "wb");
    if (!fi || !fo) { fprintf(stderr, "ref decode tap: cannot open files\n"); return 2; }
    modesChecksumInit(1);            /* what modesInit() does for the decoder (dump1090.c:206-208) */
    icaoFilterInit();
    modeACInit();
    unsigned char frame[14];
    while (fread(frame, 1, 14, fi) == 14) {
        struct modesMessage mm;
        adsbcu_fields_es e;
        int32_t rc;
        adsbcu_fields_commb b;
        memset(&mm, 0, sizeof(mm));  /* demod_2400.c:355 */
        /* Address/Parity replies are only accepted from aircraft the filter knows: let it know this one, as an earlier
         * DF11 / DF17 of the same aircraft would have (mode_s.c:781-788) */
        struct hdr_fields hf;
        {
            const int df = frame[0] >> 3;
            if (df == 20 || df == 21 || df == 16) icaoFilterAdd(modesChecksum(frame, 112));
            if (df == 0 || df == 4 || df == 5) icaoFilterAdd(modesChecksum(frame, 56));
            if (df == 11) icaoFilterAdd(((uint32_t) frame[1] << 16) | ((uint32_t) frame[2] << 8) | frame[3]);
        }
        mm.score = scoreModesMessage(frame);
        rc = decodeModesMessage(&mm, frame);
        if (rc < 0) { memset(&e, 0, sizeof(e)); memset(&b, 0, sizeof(b)); memset(&hf, 0, sizeof(hf)); }
        else { es_record(&mm, &e); mb_record(&mm, &b); hdr_record(&mm, &hf); }
        fwrite(&e, sizeof(e), 1, fo);
        fwrite(&b, sizeof(b), 1, fo);
        fwrite(&hf, sizeof(hf), 1, fo);
        fwrite(&rc, sizeof(rc), 1, fo);
    }
    fclose(fi); fclose(fo);
    return 0;
}
