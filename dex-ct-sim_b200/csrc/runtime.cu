// Error string, device checks and the version entry point of libdexct.so.
#include "common.cuh"
#include <string.h>
#include <mutex>

namespace dexct {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static int count_sm100()
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    int ok = 0;
    for (int d = 0; d < n; ++d) {
        int major = 0;
        if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, d) == cudaSuccess && major == 10) ++ok;
    }
    return ok;
}

// The library is compiled for sm_100a only and has no CPU path: refuse loudly on anything else.
int check_device()
{
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) {
        cudaGetLastError();
        set_error("no CUDA device: libdexct has no CPU fallback");
        return DEXCT_ENODEVICE;
    }
    static thread_local int cached_dev = -1, cached_ok = 0;
    if (cached_dev != dev) {
        int major = 0;
        cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
        cached_dev = dev;
        cached_ok = major == 10;
    }
    if (!cached_ok) {
        set_error("device %d is not compute capability 10.x (B200): libdexct is built for sm_100a only", dev);
        return DEXCT_ENODEVICE;
    }
    return DEXCT_OK;
}

int sm_count()
{
    static thread_local int dev_cached = -1, sms = 148;
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev != dev_cached) {
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        dev_cached = dev;
    }
    return sms;
}

namespace {
constexpr int MAX_DEV = 64;
struct SlotState {
    std::mutex mtx;
    cudaEvent_t ev[MAX_DEV] = {};
    void *staging[MAX_DEV] = {};
};
SlotState g_slots[SLOT_COUNT];
}  // namespace

ConstSlot::ConstSlot(int slot, cudaStream_t stream) : slot_(slot), dev_(0), stream_(stream), ok_(false), staging_(nullptr)
{
    SlotState &S = g_slots[slot_];
    S.mtx.lock();
    if (cudaGetDevice(&dev_) != cudaSuccess || dev_ < 0 || dev_ >= MAX_DEV) {
        set_error("ConstSlot: cannot identify the current device");
        return;
    }
    if (!S.ev[dev_]) {
        if (cudaEventCreateWithFlags(&S.ev[dev_], cudaEventDisableTiming) != cudaSuccess ||
            cudaMalloc(&S.staging[dev_], SLOT_STAGING_BYTES) != cudaSuccess) {
            set_error("ConstSlot: cannot create the table staging buffer: %s", cudaGetErrorString(cudaGetLastError()));
            return;
        }
    } else if (cudaStreamWaitEvent(stream_, S.ev[dev_], 0) != cudaSuccess) {
        set_error("ConstSlot: cudaStreamWaitEvent failed: %s", cudaGetErrorString(cudaGetLastError()));
        return;
    }
    staging_ = S.staging[dev_];
    ok_ = true;
}

ConstSlot::~ConstSlot()
{
    SlotState &S = g_slots[slot_];
    if (ok_) cudaEventRecord(S.ev[dev_], stream_);
    S.mtx.unlock();
}

}  // namespace dexct

extern "C" int dexct_version(void) { return DEXCT_VERSION; }
extern "C" const char *dexct_last_error(void) { return dexct::g_err; }
extern "C" int dexct_device_count(void) { return dexct::count_sm100(); }
