// TEST INFRASTRUCTURE ONLY (see include/cuda_runtime.h): the fiber scheduler of the SIMT emulator.
// One CTA at a time, one ucontext fiber per CUDA thread, all on the launching OS thread.
#include <cuda_runtime.h>

#include <mutex>
#include <sys/mman.h>
#include <ucontext.h>
#include <vector>

uint3 threadIdx, blockIdx;
dim3 blockDim, gridDim;

namespace emu {
namespace {

constexpr size_t kStack = 256 * 1024;
constexpr int kMaxThreads = 1024;
enum State { RUN, WAIT_CTA, WAIT_WARP, DONE };

struct Fiber {
    ucontext_t ctx;
    char *stack = nullptr;
    State st = DONE;
    int bar_id = 0, bar_count = 0;
};

Fiber g_f[kMaxThreads];
ucontext_t g_main;
int g_cur = -1, g_n = 0;
const std::function<void()> *g_body = nullptr;
alignas(128) unsigned char g_smem[232 * 1024];
unsigned long long g_slots[kMaxThreads / 32][32];
std::mutex g_launch_mutex;

void entry()
{
    (*g_body)();
    g_f[g_cur].st = DONE;
    swapcontext(&g_f[g_cur].ctx, &g_main);
}

void yield_as(State s)
{
    Fiber &f = g_f[g_cur];
    f.st = s;
    swapcontext(&f.ctx, &g_main);
}

void resume(int t)
{
    g_cur = t;
    threadIdx = uint3{unsigned(t), 0, 0};
    swapcontext(&g_main, &g_f[t].ctx);
    g_cur = -1;
}

[[noreturn]] void die(const char *what)
{
    std::fprintf(stderr, "emu: %s (block %u of %u, %d threads)\n", what, blockIdx.x, gridDim.x, g_n);
    for (int t = 0; t < g_n; ++t)
        if (g_f[t].st != DONE) { std::fprintf(stderr, "  thread %d state %d barrier %d/%d\n", t, int(g_f[t].st), g_f[t].bar_id, g_f[t].bar_count); break; }
    std::abort();
}

void run_block()
{
    for (int t = 0; t < g_n; ++t) {
        Fiber &f = g_f[t];
        if (!f.stack) {
            f.stack = static_cast<char *>(mmap(nullptr, kStack, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0));
            if (f.stack == MAP_FAILED) die("mmap of a fiber stack failed");
        }
        getcontext(&f.ctx);
        f.ctx.uc_stack.ss_sp = f.stack;
        f.ctx.uc_stack.ss_size = kStack;
        f.ctx.uc_link = nullptr;
        makecontext(&f.ctx, entry, 0);
        f.st = RUN;
    }
    const int nwarps = (g_n + 31) / 32;
    for (;;) {
        bool progress = false, all_done = true;
        for (int w = 0; w < nwarps; ++w) {
            const int lo = 32 * w, hi = std::min(g_n, lo + 32);
            for (;;) {                            // keep one warp going through its warp-level rendezvous
                bool ran = false;
                for (int t = lo; t < hi; ++t)
                    if (g_f[t].st == RUN) { resume(t); ran = true; progress = true; }
                int live = 0, waiting = 0;
                for (int t = lo; t < hi; ++t) {
                    live += g_f[t].st != DONE;
                    waiting += g_f[t].st == WAIT_WARP;
                }
                if (live && waiting == live) {
                    for (int t = lo; t < hi; ++t) if (g_f[t].st == WAIT_WARP) g_f[t].st = RUN;
                    progress = true;
                    continue;
                }
                if (waiting && !ran) break;       // some lanes sit at a CTA barrier while others wait for the warp: resolved below or deadlock
                if (!ran) break;
            }
        }
        // CTA-level barriers: release an id when `count` threads (or every live thread) wait on it
        int live = 0;
        for (int t = 0; t < g_n; ++t) live += g_f[t].st != DONE;
        all_done = live == 0;
        if (all_done) return;
        int ids[17] = {};
        int want[17] = {};
        for (int t = 0; t < g_n; ++t)
            if (g_f[t].st == WAIT_CTA) { ++ids[g_f[t].bar_id]; want[g_f[t].bar_id] = g_f[t].bar_count; }
        for (int id = 0; id < 17; ++id) {
            if (!ids[id]) continue;
            const int need = want[id] ? want[id] : live;
            if (ids[id] > need) die("more threads at a named barrier than its count");
            if (ids[id] == need) {
                for (int t = 0; t < g_n; ++t) if (g_f[t].st == WAIT_CTA && g_f[t].bar_id == id) g_f[t].st = RUN;
                progress = true;
            }
        }
        if (!progress) die("deadlock: no runnable thread and no barrier complete");
    }
}

}  // namespace

void launch(dim3 grid, dim3 block, size_t dyn_smem, const std::function<void()> &body)
{
    std::lock_guard<std::mutex> guard(g_launch_mutex);
    const unsigned n = block.x * block.y * block.z;
    if (block.y != 1 || block.z != 1 || grid.y != 1 || grid.z != 1) { std::fprintf(stderr, "emu: only 1-D launches\n"); std::abort(); }
    if (!n || n > unsigned(kMaxThreads) || dyn_smem > sizeof(g_smem)) { std::fprintf(stderr, "emu: bad launch configuration\n"); std::abort(); }
    g_body = &body;
    g_n = int(n);
    blockDim = block;
    gridDim = grid;
    for (unsigned b = 0; b < grid.x; ++b) {
        blockIdx = uint3{b, 0, 0};
        run_block();
    }
    g_body = nullptr;
}

void barrier(int id, int count)
{
    if (id < 0 || id > 16) die("barrier id out of range");
    g_f[g_cur].bar_id = id;
    g_f[g_cur].bar_count = count;
    yield_as(WAIT_CTA);
}

void warp_sync() { yield_as(WAIT_WARP); }
unsigned long long *warp_slots() { return g_slots[g_cur / 32]; }
int lane() { return g_cur & 31; }
unsigned warp_live_mask()
{
    const int lo = g_cur & ~31;
    unsigned m = 0;
    for (int k = 0; k < 32 && lo + k < g_n; ++k) if (g_f[lo + k].st != DONE) m |= 1u << k;
    return m;
}
unsigned char *dyn_smem() { return g_smem; }
int sm_count()
{
    const char *e = std::getenv("ADSBCU_EMU_SMS");
    const int n = e ? std::atoi(e) : 2;
    return n > 0 ? n : 2;
}

}  // namespace emu

// ---- "device memory" for cudaPointerGetAttributes: ranges a test marks (nbytes == 0 forgets them all) ----------------------
#include <mutex>
#include <utility>
#include <vector>
namespace {
std::mutex g_dev_mu;
std::vector<std::pair<const char *, size_t>> g_dev_ranges;
}  // namespace
extern "C" void adsbcu_emu_mark_device(const void *p, size_t nbytes)
{
    std::lock_guard<std::mutex> lock(g_dev_mu);
    if (!nbytes) g_dev_ranges.clear();
    else g_dev_ranges.emplace_back(static_cast<const char *>(p), nbytes);
}
extern "C" int adsbcu_emu_is_device(const void *p)
{
    std::lock_guard<std::mutex> lock(g_dev_mu);
    for (const auto &r : g_dev_ranges)
        if (static_cast<const char *>(p) >= r.first && static_cast<const char *>(p) < r.first + r.second) return 1;
    return 0;
}
