// Shared between the host side (adsbcu_api.cpp, adsbcu_tables.cpp) and the
// kernels (adsbcu_front.cu, adsbcu_msgs.cu, adsbcu_finish.cu, adsbcu_kernels.cu).  Domain vocabulary follows the reference:
// samples, magnitudes, preambles, phases, syndromes, the ICAO filter.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

// Kernel launches and dynamic shared memory go through two macros so that the CPU test suite can run these very
// sources under its SIMT emulator (tests/emu, test infrastructure only: g++ with a shadow <cuda_runtime.h> that
// defines ADSBCU_EMU).  The product is always the nvcc build for sm_100a, where they are the plain CUDA forms.
#ifdef ADSBCU_EMU
#define ADSBCU_LAUNCH(kernel, grid, block, smem, stream, ...) emu::launch(dim3(grid), dim3(block), (smem), [&] { kernel(__VA_ARGS__); })
#define ADSBCU_DYN_SMEM(name) unsigned char *name = emu::dyn_smem()
#else
#define ADSBCU_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#define ADSBCU_DYN_SMEM(name) extern __shared__ __align__(16) unsigned char name[]
#endif

struct adsbcu_msg;
namespace adsbcu {

constexpr int      kOverlap       = 326;      // dump1090.c:159 trailing_samples
constexpr int      kLookahead     = 290;      // samples read past a start position (demod_2400.c:121, +1 pad)
constexpr uint32_t kNtFlag        = 1u << 25; // icao_filter.h:24
constexpr uint32_t kNonIcao       = 1u << 24; // dump1090.h:262
constexpr uint32_t kUnchecked     = 0xFFFFFFFFu;

// State-independent classification of one sliced phase.  The host combines it
// with the filter state to get the reference's score_rank (mode_s.h:32-80).
enum Kind : uint32_t {
    K_AP = 0,          // DF0/4/5/16/20/21: address = syndrome; accepted only if known
    K_DF11_ACQ_0,      // DF11, IID 0, clean
    K_DF11_IID_0,      // DF11, IID != 0, clean
    K_DF11_ACQ_1,      // DF11, IID 0, one bit repaired
    K_DF11_IID_1,
    K_DF17_0, K_DF17_1, K_DF17_2,
    K_DF18_0, K_DF18_1, K_DF18_2,
    K_COUNT
};

// rank when the address is unknown / known
__host__ __device__ inline int rank_of(uint32_t kind, bool known)
{
    switch (kind) {
    case K_AP:         return known ? 12 : 5;
    case K_DF11_ACQ_0: return known ? 23 : 17;
    case K_DF11_IID_0: return known ? 20 : 8;
    case K_DF11_ACQ_1: return known ? 19 : 7;
    case K_DF11_IID_1: return known ? 18 : 6;
    case K_DF17_0:     return known ? 27 : 25;
    case K_DF17_1:     return known ? 22 : 16;
    case K_DF17_2:     return known ? 14 : 10;
    case K_DF18_0:     return known ? 26 : 24;
    case K_DF18_1:     return known ? 21 : 15;
    case K_DF18_2:     return known ? 13 : 9;
    default:           return 0;
    }
}
constexpr int kAcceptThreshold = 11;          // SR_ACCEPT_THRESHOLD

// (position, phase) record kept by the scan kernel: 8 bytes.
//   pos  : start position relative to the feed (padded-stream coordinate minus feed base)
//   info : bits 0..24 filter key (24-bit address, bit 24 = DF18 "NT" table), 25..27 phase-4, 28..31 Kind
struct PhaseRec {
    uint32_t pos;
    uint32_t info;
};

// Fully evaluated record produced by the finalize kernel for the host pass.
struct alignas(8) FullRec {
    uint32_t pos;          // as PhaseRec.pos
    uint8_t  phase;        // try_phase 4..8
    uint8_t  kind;         // Kind
    int8_t   corrections;  // 0..2
    uint8_t  msgtype;      // DF after correction
    uint8_t  verbatim[14];
    uint8_t  msg[14];
    uint32_t crc;          // mm.crc
    uint32_t addr;         // mm.addr (with non-ICAO flag for DF18)
    uint32_t test_key;     // key for icaoFilterTest
    uint32_t add_key;      // key for icaoFilterAdd, kUnchecked if the message is not added
    double   signal_level; // mean squared magnitude over the signal window, full scale = 1.0
};
static_assert(sizeof(FullRec) == 64, "FullRec layout");

// Syndrome hash tables (open addressing, linear probe, power-of-two slots).
// entry: bits 0..23 syndrome, 24..25 nerr, 26..32 bit0, 33..39 bit1; all-ones = empty.
constexpr int      kSyndSlotsLong  = 8192;
constexpr int      kSyndSlotsShort = 4096;
constexpr uint64_t kSyndEmpty      = ~0ull;
__host__ __device__ inline uint32_t synd_hash(uint32_t s, uint32_t slots) { return (s * 2654435761u) >> 8 & (slots - 1); }

struct DevTables {
    float    sq[256];            // ((i-127.4)/128)^2 as float, scaled by 2^34 (tables.c:24-29)
    uint32_t bit_synd[112];      // CRC-24 syndrome of each single bit of a 112-bit frame (crc.c:57-62)
    uint32_t valid_short, valid_long;       // DF pre-filter (demod_2400.c:80-93)
    uint32_t fixable_short, fixable_long;   // df_correctable_* for the active setting (mode_s.c:260-265)
    int32_t  nfix;
    int32_t  pad;
    uint64_t synd_long[kSyndSlotsLong];
    uint64_t synd_short[kSyndSlotsShort];
    alignas(16) uint32_t occ_long[kSyndSlotsLong / 32];    // bit h: synd_long[h] is occupied (the scan keeps these in shared
    alignas(16) uint32_t occ_short[kSyndSlotsShort / 32];  // memory: most probes end at an empty slot without a global load)
    alignas(16) uint32_t crc_byte[11][256];  // syndrome contribution of byte b (0..10) of a 112-bit frame having value v
    alignas(16) uint16_t mag_lut[32768];   // central 128 x 128 part (I, Q in [64, 192)) of the reference's uc8 magnitude
                                 // table (tables.c:9-41) in the bank-swizzled layout of lut_offset(); filled on the
                                 // device by build_lut at create time.  Samples outside it (|I-128| > 64: strong
                                 // signals only) take the register path mag_iq().
    alignas(16) uint16_t mag_lut_full[65536];   // the whole table, element lutf_index(I | Q << 8): front kernel (128 KiB of shared memory)
};

// Byte offset of magnitude(I, Q) inside mag_lut, for I, Q in [64, 192): [64,192) -> (x & 127) is a bijection
// onto [0,128).  Element index = (Q & 127) << 8 | (I & 127): the two 7-bit fields stay where they are in the
// 16-bit sample word, so one AND addresses both samples of a 32-bit word at once (bit 7 of the index is a hole:
// rows are 256 elements of which 128 are used).  Five low bits of Q are xor-ed into index bits 1..5 (= the
// shared-memory bank) so that the 32 lanes of a gather spread over the banks even when every sample sits
// within a few LSB of the DC point.
__host__ __device__ inline uint32_t lut_offset(uint32_t i, uint32_t q)
{
    return 2u * ((((q & 127u) << 8) | (i & 127u)) ^ ((q & 31u) << 1));
}
__host__ __device__ inline bool lut_covers(uint32_t x) { return x >= 64u && x < 192u; }

// Per-chunk counters (device; copied to pinned host memory after the chunk's kernels).
struct ChunkCounters {
    unsigned long long preambles;
    unsigned int n_cand;         // (position, phase) records written by the scan kernel (all segments)
    unsigned int n_full;         // FullRec written by the evaluate kernel
    unsigned int overflow;       // a tile list, a segment or full[] overflowed: results are incomplete
    unsigned int pad;
    unsigned int n_msgs;         // messages the ordered pass accepted in this chunk
    unsigned int n_sequential;   // records the ordered pass took through its one-thread replay (0 in the steady state)
    unsigned long long msg_base; // ... and how many the feed had accepted before it
    unsigned long long dbg[8];   // -DADSBCU_ORDER_TIMING: globaltimer marks of the order kernel
};

// One chunk = a contiguous range of start positions of the current feed.  Position p (feed-relative,
// in the reference's padded mag_buf coordinate) looks at feed samples p-326 .. p-326+289; samples
// before the feed come from carry_mag (fifo.c:187-196), so a chunk never reads past the feed end.
//
// Ordered emission: segment g of `nseg` owns a contiguous run of the chunk's tiles and writes its
// records in (position, phase) order into seg[g*seg_cap ..]; segments in index order are therefore
// the whole chunk in stream order and the host pass never sorts.
struct ScanArgs {
    const uint8_t  *iq;          // feed samples (device), 16 B aligned
    const uint16_t *mag_in;      // or, for SC16 / SC16Q11 inputs, their magnitudes (then iq is unused); 16 B aligned
    const uint16_t *carry_mag;   // kOverlap magnitudes preceding the feed (zeros at stream start)
    uint32_t        nsamples;    // samples in the feed
    uint32_t        pos_begin, pos_end;   // this chunk's start positions
    const DevTables *tab;
    uint32_t       *bitmap;      // 2 x 2^24 bits: [plain | NT] addresses that may enter the filter
    PhaseRec       *seg;         uint32_t seg_cap;  uint32_t nseg;
    uint32_t       *seg_count;   // [nseg] records per segment (scan: kept; filter: survivors)
    FullRec        *full;        uint32_t full_cap; // mapped pinned host memory
    ChunkCounters  *ctr;
    // sharded captures only (else null): records the filter drops ONLY for an address missing from the
    // bitmap are appended here (unordered); another shard's addresses may revive them later
    PhaseRec       *deferred;    uint32_t deferred_cap;
    uint32_t       *n_deferred;
};

void build_tables(DevTables &t, int nfix_crc, int fix_df);

// ---- round-2 pipeline: front kernel -> message kernel -> finish kernel --------------------------------------
// front kernel geometry: kFWarps autonomous warps per CTA, a lane owns kFRun consecutive samples of a warp-tile
#ifndef ADSBCU_FWARPS                 // warps per front-kernel CTA (measurement knob; fewer leave registers for message-kernel CTAs on the same SM)
#define ADSBCU_FWARPS 32
#endif
constexpr int kFRun = 24, kFWarps = ADSBCU_FWARPS;
constexpr int kWTile = kFRun * 32;            // 768 samples staged per warp-tile == start positions per warp-tile
constexpr int kFLead = 2;                     // staged index of a tile's first start position (16 B alignment of the I/Q vectors)
constexpr int kFHalo = 24;                    // magnitudes staged past the tile: the preamble test looks 18 samples ahead
constexpr int kFNeed = 296;                   // samples past a chunk's last start position whose slicer decisions its messages use
                                              // (lead 2 + 19 + 269 data samples + 3 correlator taps, rounded up)
// element of mag_lut_full for the 16-bit sample s = I | Q << 8: three low bits of Q are xor-ed into index bits 3..5.
// The shared-memory bank is index bits 1..5 and bit 0 picks the half of the bank word, so an 8 x 8 window of (I, Q)
// around the DC point maps onto 64 different (bank, half) slots: the gather of near-DC noise is almost conflict-free
__host__ __device__ inline uint32_t lutf_index(uint32_t s) { return s ^ ((s >> 5) & 0x38u); }

// One chunk of a feed.  Position p (feed-relative, the reference's padded mag_buf coordinate) looks at feed samples
// p-326 .. p-326+289.  Segment g of `nseg` (one front-kernel CTA) owns a contiguous run of the chunk's warp-tiles;
// segments in index order are the chunk in stream order.
struct ChunkArgs {
    const uint8_t  *iq;          // feed samples (device), 16 B aligned
    const uint16_t *mag_in;      // or, for SC16 / SC16Q11 inputs, their magnitudes (then iq is unused); 16 B aligned
    const uint16_t *carry_mag;   // kOverlap magnitudes preceding the feed (zeros at stream start)
    uint32_t        nsamples;    // feed samples this chunk may read (a host feed is uploaded chunk by chunk)
    uint32_t        pos_begin, pos_end;   // this chunk's start positions; pos_begin is a multiple of 8
    const DevTables *tab;
    uint32_t       *bitmap;      // 2 x 2^24 bits: [plain | NT] addresses that may enter the filter
    uint32_t        nseg;
    uint16_t       *planes;      uint32_t plane_stride;   // [12][plane_stride] 10-bit fields, field index = chunk-local (sample + kFLead) / 24
    uint32_t       *pre;         uint32_t pre_cap;   uint32_t *pre_count;    // [nseg][pre_cap] confirmed preambles, chunk-local positions
    PhaseRec       *pend;        uint32_t pend_cap;  uint32_t *pend_count;   // [nseg][pend_cap] parked records (address not announced yet)
    FullRec        *rec;         uint32_t rec_cap;   uint32_t *rec_count;    // [nseg][rec_cap] evaluated records, unordered inside a segment
    FullRec        *ordered;     uint32_t full_cap;  // the chunk's records in stream order (device memory)
    ChunkCounters  *ctr;
};

// The reference's ICAO filter (icao_filter.c:23-109) and demodulate2400's carried skip state, device resident.
constexpr uint32_t kIcaoSlots = 4096, kIcaoEmpty = 0xFFFFFFFFu;
struct StreamState {
    uint32_t table[2][kIcaoSlots];   // icao_filter_a / icao_filter_b
    uint32_t active;                 // index of icao_filter_active
    uint32_t active_count;           // occupied slots of the active table
    uint32_t has_last;               // last accepted message: its start and last_message_end, both in
    uint32_t poison;                 // a chunk of the running feed overflowed: later chunks must leave the state alone
    uint64_t last_a, last_lme;       // stream coordinates counted from the first (zero) overlap sample
    uint64_t n_out;                  // messages accepted so far in the running feed
};

struct FinishArgs {
    ChunkArgs    c;
    StreamState *st;
    struct ::adsbcu_msg *out;    uint64_t out_cap;       // accepted messages of the feed, stream order (device-visible memory)
    uint2       *accepted;       // [full_cap] (index into ordered[], score_rank) of this chunk's accepted messages, stream order
    uint64_t     feed_base;      // stream coordinate of the feed's position 0
    uint32_t     block_samples;  // the reference's mag_buf length (skip rule at block boundaries, timestamps)
    uint32_t     candidates_only;    // sharded scans: stop after the ordered array
    uint32_t     force_sequential;   // test knob: always take the one-thread replay
    uint32_t     batch;              // records per batch of the ordered pass (test knob; the kernel caps it at its shared-memory size)
    unsigned long long *seg_flag;    // [nseg] epoch << 32 | records of the segment, published by its CTA
    uint32_t     epoch;          // chunk sequence number (never 0): tells this chunk's flags from older ones
    uint32_t    *run_start;      // [full_cap + 1]
    uint4       *view;           // [full_cap] 16-byte view of the ordered records for the ordered pass
    PhaseRec    *deferred;       uint32_t deferred_cap;  uint32_t *n_deferred;   // sharded scans: still-unknown parked records
};

// launchers of the round-2 pipeline
cudaError_t init_front();
cudaError_t init_finish();
cudaError_t launch_build_lutf(DevTables *tab, cudaStream_t s);
cudaError_t init_messages();
cudaError_t launch_front(const ChunkArgs &a, cudaStream_t s);        // grid = a.nseg
cudaError_t launch_messages(const ChunkArgs &a, uint32_t grid, cudaStream_t s);
cudaError_t launch_order(const FinishArgs &fa, cudaStream_t s);      // late records, stream order (one CTA per segment)
cudaError_t launch_resolve(const FinishArgs &fa, cudaStream_t s);    // the ordered pass (one CTA)
cudaError_t launch_level(const FinishArgs &fa, uint32_t level_grid, cudaStream_t s);   // signal levels, final messages
cudaError_t launch_state_reset(StreamState *st, int clear_filter, int clear_stream, cudaStream_t s);
cudaError_t launch_state_begin_feed(StreamState *st, int reset_count, cudaStream_t s);
cudaError_t launch_filter_add(StreamState *st, uint32_t *bitmap, uint32_t addr, cudaStream_t s);
cudaError_t launch_filter_test(const StreamState *st, uint32_t addr, uint32_t *result_dev, cudaStream_t s);
cudaError_t launch_filter_expire(StreamState *st, uint32_t *bitmap, uint32_t nwords, cudaStream_t s);
cudaError_t launch_state_import(StreamState *st, uint32_t *bitmap, uint32_t has_last, uint64_t last_a, uint64_t last_lme,
                                const uint32_t *addrs_dev, uint32_t n, cudaStream_t s);

// launchers (adsbcu_kernels.cu)
cudaError_t init_kernels();                                        // per device, once (opt-in shared memory size)
// deferred records whose address is in the bitmap now -> flat into a.seg (seg_count set per segment)
cudaError_t launch_revive(const ScanArgs &a, cudaStream_t s);
cudaError_t launch_or_words(uint32_t *dst, const uint32_t *src, uint32_t nwords, cudaStream_t s);
cudaError_t launch_build_lut(DevTables *tab, cudaStream_t s);     // fills tab->mag_lut with the register-computed magnitudes
cudaError_t launch_carry(const uint8_t *iq, uint32_t nsamples, uint16_t *carry_mag, const DevTables *tab, cudaStream_t s);
cudaError_t launch_convert_sc16(const int16_t *in, uint32_t nsamples, uint16_t *mag, int q11, cudaStream_t s);
cudaError_t launch_carry_mag(const uint16_t *mag_in, uint32_t nsamples, uint16_t *carry_mag, cudaStream_t s);
cudaError_t launch_magnitude(const uint8_t *iq, uint32_t nsamples, uint16_t *mag, const DevTables *tab, cudaStream_t s);


// Mode A/C (adsbcu_modeac.cu): demodulate2400AC on one mag_buf
struct AcRec { uint32_t f1, f2_clock, modeac; };
constexpr int kAcLead = 2;    // magnitudes buffer: m[0] at base + kAcLead puts the first new sample m[326] on 16 bytes
cudaError_t launch_modeac(const DevTables *tab, int sm_count, cudaStream_t s, const uint8_t *d_iq_block, uint32_t n,
                          const uint8_t *d_iq_overlap, uint32_t n_overlap, uint16_t *d_mag_base, unsigned long long *d_sums,
                          AcRec *d_rec, uint32_t cap, uint32_t *d_count);

// SM partition (adsbcu_partition.cpp): streams bound to disjoint SM sets, dense kernels / post kernels
struct SmPartition {
    bool ok = false;
    int sm_main = 0, sm_post = 0;
    cudaStream_t main = nullptr, msg = nullptr, order = nullptr, resolve = nullptr, level = nullptr;
};
bool make_partition(int device, int post_sms, SmPartition &p);
void free_partition(SmPartition &p);

// Path B (adsbcu_pathb.cu): libadsb's own 2.0 MS/s demodulator
struct PathB;
}  // namespace adsbcu
#include <string>
struct adsbcu_handle;
namespace adsbcu {
// What an entry point that lives outside adsbcu_api.cpp may know about a handle (adsbcu_api.cpp defines both).
struct HandleView { int device, sm_count; const DevTables *tab; bool host_only; };
HandleView handle_view(const adsbcu_handle *h);
void handle_set_error(adsbcu_handle *h, const char *what);
}  // namespace adsbcu
struct adsbcu_msg_b;
namespace adsbcu {
int  pathb_feed(PathB *&b, const DevTables *d_tab, int sm_count, cudaStream_t s, const uint8_t *iq, size_t nbytes,
                adsbcu_msg_b *out, size_t cap, size_t *n_out, std::string &err);
void pathb_reset(PathB *b);
void pathb_free(PathB *b);
}  // namespace adsbcu
struct adsbcu_uat_frame;
struct adsbcu_fields_b;
struct adsbcu_cpr_pair;
struct adsbcu_cpr_fix;
struct adsbcu_cpr_in;
struct adsbcu_cpr_out;
struct adsbcu_fields_a;
struct adsbcu_fields_es;
struct adsbcu_fields_commb;
namespace adsbcu {
// Path B, next layer (adsbcu_fields.cu): bulk field decode and CPR
struct Fields;
int  fields_decode(Fields *&f, cudaStream_t s, const adsbcu_msg_b *msgs, size_t n, adsbcu_fields_b *out, std::string &err);
int  cpr_decode(Fields *&f, cudaStream_t s, const adsbcu_cpr_pair *pairs, size_t n, adsbcu_cpr_fix *out, std::string &err);
int  cpr_a_decode(Fields *&f, cudaStream_t s, const adsbcu_cpr_in *in, size_t n, adsbcu_cpr_out *out, std::string &err);
int  fields_a_decode(Fields *&f, cudaStream_t s, const ::adsbcu_msg *msgs, size_t n, ::adsbcu_fields_a *out, std::string &err);
int  fields_es_decode(Fields *&f, cudaStream_t s, const ::adsbcu_msg *msgs, size_t n, ::adsbcu_fields_es *out, std::string &err);
int  fields_commb_decode(Fields *&f, cudaStream_t s, const ::adsbcu_msg *msgs, size_t n, ::adsbcu_fields_commb *out, std::string &err);
int  fields_all_device(int sm_count, cudaStream_t s, const ::adsbcu_msg *msgs_dev, size_t n, ::adsbcu_fields_a *a_dev, ::adsbcu_fields_es *es_dev,
                       ::adsbcu_fields_commb *mb_dev, std::string &err);
void fields_free(Fields *f);

// Path C (adsbcu_uat.cu): 978 MHz UAT
struct PathC;
int  pathc_feed(PathC *&c, int sm_count, cudaStream_t s, const uint8_t *iq, size_t nbytes, bool on_device, adsbcu_uat_frame *out, size_t cap,
                size_t *n_out, std::string &err);
void pathc_stats(const PathC *c, float *ms_device, uint32_t *calls, uint32_t *launches);
int  pathc_phase(PathC *&c, int sm_count, cudaStream_t s, const uint8_t *iq, size_t n, uint16_t *phi_out, std::string &err);
void pathc_reset(PathC *c);
void pathc_free(PathC *c);

}  // namespace adsbcu
