// SM partition for the post kernels (order / resolve / level).
//
// The ordered pass of a chunk is a chain resolve(c-1) -> resolve(c) that runs beside the dense kernels of later chunks.
// A resolve CTA needs a whole SM (1024 threads, ~220 KiB of shared memory); on a GPU whose SMs are all taken by
// front / message CTAs it waits for one to drain (~50 us per chunk, measured with profiles/order_probe.py), and that
// wait is on the critical path of the feed.  CUDA green contexts (driver API, CUDA >= 12.4) give streams that are
// bound to a subset of the SMs: a few SMs (the minimum split on sm_90+ is 8) are set aside for the three post streams,
// the dense kernels get the rest.  The driver entry points are fetched with cudaGetDriverEntryPoint, so the library
// has no link-time dependency on libcuda; if anything here fails the engine keeps its single-partition scheme.
#include "adsbcu_internal.h"

#ifndef ADSBCU_EMU
#include <cuda.h>
#include <cudaTypedefs.h>
#endif

#include <cstdio>
#include <cstdlib>
#include <mutex>

namespace adsbcu {

#ifdef ADSBCU_EMU
bool make_partition(int, int, SmPartition &p) { p = SmPartition{}; return false; }
void free_partition(SmPartition &p) { p = SmPartition{}; }
#else

namespace {
template <class F>
bool entry(const char *name, F &fn)
{
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q{};
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
        cudaGetLastError();
        return false;
    }
    fn = reinterpret_cast<F>(p);
    return true;
}
}  // namespace

bool make_partition(int device, int post_sms, SmPartition &p)
{
    p = SmPartition{};
    PFN_cuDeviceGet_v2000 f_dev = nullptr;
    PFN_cuDeviceGetDevResource_v12040 f_res = nullptr;
    PFN_cuDevSmResourceSplitByCount_v12040 f_split = nullptr;
    PFN_cuDevResourceGenerateDesc_v12040 f_desc = nullptr;
    PFN_cuGreenCtxCreate_v12040 f_create = nullptr;
    PFN_cuGreenCtxStreamCreate_v12050 f_stream = nullptr;
    if (!entry("cuDeviceGet", f_dev) || !entry("cuDeviceGetDevResource", f_res) || !entry("cuDevSmResourceSplitByCount", f_split) ||
        !entry("cuDevResourceGenerateDesc", f_desc) || !entry("cuGreenCtxCreate", f_create) || !entry("cuGreenCtxStreamCreate", f_stream))
        return false;
    // the two green contexts of a device are made once per process and shared by all handles; every handle gets its own streams
    struct Cached { int state = 0; int post_sms = 0; CUgreenCtx g_main = nullptr, g_post = nullptr; int sm_main = 0, sm_post = 0; };
    static std::mutex mu;
    static Cached cache[64];
    if (device < 0 || device >= 64) return false;
    std::lock_guard<std::mutex> lock(mu);
    Cached &c = cache[device];
    if (c.state == 0 || (c.state == 1 && c.post_sms != post_sms)) {
        c = Cached{};
        c.state = -1;                                        // failed, unless everything below works
        c.post_sms = post_sms;
        CUdevice dev;
        if (f_dev(&dev, device) != CUDA_SUCCESS) return false;
        CUdevResource all{}, post{}, rest{};
        if (f_res(dev, &all, CU_DEV_RESOURCE_TYPE_SM) != CUDA_SUCCESS) return false;
        unsigned groups = 1;
        if (f_split(&post, &groups, &all, &rest, 0, unsigned(post_sms)) != CUDA_SUCCESS || groups != 1) return false;
        if (post.sm.smCount == 0 || rest.sm.smCount < 64) return false;
        CUdevResourceDesc d_post = nullptr, d_main = nullptr;
        if (f_desc(&d_post, &post, 1) != CUDA_SUCCESS || f_desc(&d_main, &rest, 1) != CUDA_SUCCESS) return false;
        if (f_create(&c.g_post, d_post, dev, CU_GREEN_CTX_DEFAULT_STREAM) != CUDA_SUCCESS) return false;
        if (f_create(&c.g_main, d_main, dev, CU_GREEN_CTX_DEFAULT_STREAM) != CUDA_SUCCESS) return false;
        c.sm_main = int(rest.sm.smCount);
        c.sm_post = int(post.sm.smCount);
        c.state = 1;
    }
    if (c.state != 1) return false;
    CUgreenCtx g_main = c.g_main, g_post = c.g_post;
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    CUstream s[5] = {};
    if (f_stream(&s[0], g_main, CU_STREAM_NON_BLOCKING, lo) != CUDA_SUCCESS || f_stream(&s[1], g_main, CU_STREAM_NON_BLOCKING, lo) != CUDA_SUCCESS ||
        f_stream(&s[2], g_post, CU_STREAM_NON_BLOCKING, hi) != CUDA_SUCCESS || f_stream(&s[3], g_post, CU_STREAM_NON_BLOCKING, hi) != CUDA_SUCCESS ||
        f_stream(&s[4], g_post, CU_STREAM_NON_BLOCKING, hi) != CUDA_SUCCESS)
        return false;
    p.ok = true;
    p.sm_main = c.sm_main;
    p.sm_post = c.sm_post;
    p.main = s[0]; p.msg = s[1]; p.order = s[2]; p.resolve = s[3]; p.level = s[4];
    if (std::getenv("ADSBCU_DEBUG_CHUNKS")) std::fprintf(stderr, "SM partition: %d SMs dense kernels, %d SMs post kernels\n", p.sm_main, p.sm_post);
    return true;
}

void free_partition(SmPartition &p) { p = SmPartition{}; }   // the streams are destroyed with the handle's other streams
#endif

}  // namespace adsbcu
