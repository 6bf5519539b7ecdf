// Peer-memory plumbing for view-sharded scans on one NVSwitch node: every rank maps the
// sinogram buffers of all others (CUDA IPC) and results are STORED straight into the peers' copies over NVLink -- by the
// decomposition kernel's own epilogue (gn.cu) and by the push kernel below -- instead of being all-gathered afterwards.
// At C2 size a step is 0.76 ms, and the fixed latency of four NCCL calls per step cost 36 % of the 8-GPU throughput
// (profiles/r01h_*); a store needs no rendezvous.  Synchronisation: epoch flags in peer memory, written with
// system-scope release stores after a __threadfence_system and polled with system-scope acquire loads.
#include "common.cuh"
#include <stdlib.h>
#include <string.h>

namespace dexct {
namespace {

constexpr int MAXP = DEXCT_MAX_PEERS;
struct PeerPtrs { void *p[MAXP]; };

__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// Spin until *flag >= epoch; gives up after `timeout_ns` (a peer died or the call order diverged) and reports it.
__device__ bool wait_flag(const unsigned long long *flag, unsigned long long epoch, unsigned long long timeout_ns)
{
    const unsigned long long t0 = globaltimer_ns();
    while (ld_acquire_sys(flag) < epoch) {
        if (globaltimer_ns() - t0 > timeout_ns) return false;
        __nanosleep(64);
    }
    return true;
}

// Control block of one rank (device memory, mapped by every peer), all 64-bit words:
//   [0, MAXP)            barrier flags: word r = last epoch rank r has arrived at
//   [MAXP, 3 MAXP)       max-exchange flags, two parities
//   [3 MAXP, 5 MAXP)     max-exchange keys, two parities
//   [5 MAXP]             status: 0 ok, otherwise 1 + the rank that was waited for in vain
constexpr int CTRL_BAR = 0, CTRL_MAXFLAG = MAXP, CTRL_MAXKEY = 3 * MAXP, CTRL_STATUS = 5 * MAXP, CTRL_WORDS = 5 * MAXP + 1;

__global__ void peer_barrier_kernel(PeerPtrs ctrl, int rank, int ws, unsigned long long epoch, unsigned long long timeout_ns)
{
    const int t = threadIdx.x;
    if (t >= ws) return;
    __threadfence_system();                                   // everything this device stored before is visible first
    st_release_sys((unsigned long long *)ctrl.p[t] + CTRL_BAR + rank, epoch);
    unsigned long long *mine = (unsigned long long *)ctrl.p[rank];
    if (!wait_flag(mine + CTRL_BAR + t, epoch, timeout_ns)) mine[CTRL_STATUS] = 1 + t;
}

// All-reduce (max) of one 64-bit integer key per rank: *key (local) <- max over ranks.
__global__ void peer_max_kernel(PeerPtrs ctrl, int rank, int ws, unsigned long long epoch, long long *key, int key_is_32,
                                unsigned long long timeout_ns)
{
    __shared__ long long sh[MAXP];
    const int t = threadIdx.x;
    const int par = (int)(epoch & 1);
    if (t < ws) {
        const long long mykey = key_is_32 ? (long long)*(const int *)key : *key;
        unsigned long long *peer = (unsigned long long *)ctrl.p[t];
        ((volatile long long *)peer)[CTRL_MAXKEY + par * MAXP + rank] = mykey;
        __threadfence_system();
        st_release_sys(peer + CTRL_MAXFLAG + par * MAXP + rank, epoch);
        unsigned long long *mine = (unsigned long long *)ctrl.p[rank];
        if (wait_flag(mine + CTRL_MAXFLAG + par * MAXP + t, epoch, timeout_ns)) {
            sh[t] = ((volatile long long *)mine)[CTRL_MAXKEY + par * MAXP + t];
        } else {
            mine[CTRL_STATUS] = 1 + t;
            sh[t] = mykey;
        }
    }
    __syncthreads();
    if (t == 0) {
        long long m = sh[0];
        for (int r = 1; r < ws; ++r) m = sh[r] > m ? sh[r] : m;
        if (key_is_32) *(int *)key = (int)m; else *key = m;
    }
}

// src[0, bytes) -> dst.p[d][0, bytes) for every d != skip; 16-byte accesses (bytes and all pointers 16-byte aligned).
__global__ void __launch_bounds__(512) peer_push_kernel(const uint4 *__restrict__ src, long long n16, PeerPtrs dst, int n_dst, int skip)
{
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
        const uint4 v = src[i];
        for (int d = 0; d < n_dst; ++d)
            if (d != skip) ((uint4 *)dst.p[d])[i] = v;
    }
}

int load_ptrs(PeerPtrs &pp, void *const *ptrs_host, int n)
{
    DEXCT_REQUIRE(ptrs_host && n >= 1 && n <= MAXP, "1 <= number of peers <= DEXCT_MAX_PEERS and a pointer table required");
    for (int i = 0; i < MAXP; ++i) pp.p[i] = i < n ? ptrs_host[i] : nullptr;
    for (int i = 0; i < n; ++i) DEXCT_REQUIRE(pp.p[i], "peer pointer %d is null", i);
    return DEXCT_OK;
}

}  // namespace
}  // namespace dexct

using namespace dexct;

extern "C" size_t dexct_peer_ctrl_bytes(void) { return sizeof(unsigned long long) * CTRL_WORDS; }

extern "C" int dexct_peer_alloc(size_t bytes, void **ptr_host)
{
    if (int e = check_device()) return e;
    DEXCT_REQUIRE(ptr_host && bytes > 0, "ptr_host must be non-null and bytes > 0");
    DEXCT_CUDA_OK(cudaMalloc(ptr_host, bytes));
    DEXCT_CUDA_OK(cudaMemset(*ptr_host, 0, bytes));
    DEXCT_CUDA_OK(cudaDeviceSynchronize());
    return DEXCT_OK;
}

extern "C" int dexct_peer_free(void *ptr)
{
    if (ptr) DEXCT_CUDA_OK(cudaFree(ptr));
    return DEXCT_OK;
}

extern "C" int dexct_peer_export(const void *ptr, void *handle64_host)
{
    if (int e = check_device()) return e;
    DEXCT_REQUIRE(ptr && handle64_host, "ptr and handle64_host must be non-null");
    static_assert(sizeof(cudaIpcMemHandle_t) == DEXCT_PEER_HANDLE_BYTES, "handle size");
    cudaIpcMemHandle_t h;
    DEXCT_CUDA_OK(cudaIpcGetMemHandle(&h, const_cast<void *>(ptr)));
    memcpy(handle64_host, &h, sizeof(h));
    return DEXCT_OK;
}

extern "C" int dexct_peer_open(const void *handle64_host, void **ptr_host)
{
    if (int e = check_device()) return e;
    DEXCT_REQUIRE(handle64_host && ptr_host, "handle64_host and ptr_host must be non-null");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64_host, sizeof(h));
    DEXCT_CUDA_OK(cudaIpcOpenMemHandle(ptr_host, h, cudaIpcMemLazyEnablePeerAccess));
    return DEXCT_OK;
}

extern "C" int dexct_peer_close(void *ptr)
{
    if (ptr) DEXCT_CUDA_OK(cudaIpcCloseMemHandle(ptr));
    return DEXCT_OK;
}

extern "C" int dexct_peer_barrier(void *const *ctrl_ptrs_host, int32_t rank, int32_t world_size, uint64_t epoch,
                                  double timeout_s, void *stream)
{
    if (int e = check_device()) return e;
    PeerPtrs pp;
    if (int e = load_ptrs(pp, ctrl_ptrs_host, world_size)) return e;
    DEXCT_REQUIRE(rank >= 0 && rank < world_size && epoch > 0 && timeout_s > 0, "0 <= rank < world_size, epoch > 0, timeout_s > 0");
    peer_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(pp, rank, world_size, epoch, (unsigned long long)(timeout_s * 1e9));
    DEXCT_CUDA_OK(cudaGetLastError());
    return DEXCT_OK;
}

extern "C" int dexct_peer_max_key(void *const *ctrl_ptrs_host, int32_t rank, int32_t world_size, uint64_t epoch, void *key,
                                  int32_t dtype, double timeout_s, void *stream)
{
    if (int e = check_device()) return e;
    PeerPtrs pp;
    if (int e = load_ptrs(pp, ctrl_ptrs_host, world_size)) return e;
    DEXCT_REQUIRE(rank >= 0 && rank < world_size && epoch > 0 && timeout_s > 0 && key, "0 <= rank < world_size, epoch > 0, timeout_s > 0, key non-null");
    DEXCT_REQUIRE(dtype == DEXCT_F32 || dtype == DEXCT_F64, "dtype must be DEXCT_F32 or DEXCT_F64");
    peer_max_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(pp, rank, world_size, epoch, (long long *)key, dtype == DEXCT_F32,
                                                       (unsigned long long)(timeout_s * 1e9));
    DEXCT_CUDA_OK(cudaGetLastError());
    return DEXCT_OK;
}

extern "C" int dexct_peer_status(const void *ctrl_local, int32_t *status_host, void *stream)
{
    if (int e = check_device()) return e;
    DEXCT_REQUIRE(ctrl_local && status_host, "ctrl_local and status_host must be non-null");
    unsigned long long s = 0;
    DEXCT_CUDA_OK(cudaMemcpyAsync(&s, (const unsigned long long *)ctrl_local + CTRL_STATUS, sizeof(s), cudaMemcpyDeviceToHost,
                                  (cudaStream_t)stream));
    DEXCT_CUDA_OK(cudaStreamSynchronize((cudaStream_t)stream));
    *status_host = (int32_t)s;
    return DEXCT_OK;
}

extern "C" int dexct_peer_push(const void *src, size_t bytes, void *const *dst_ptrs_host, int32_t n_dst, int32_t skip,
                               int32_t max_ctas, void *stream)
{
    if (int e = check_device()) return e;
    if (bytes == 0) return DEXCT_OK;
    PeerPtrs pp;
    if (int e = load_ptrs(pp, dst_ptrs_host, n_dst)) return e;
    DEXCT_REQUIRE(src && bytes % 16 == 0 && ((uintptr_t)src & 15) == 0, "src must be non-null, 16-byte aligned, bytes a multiple of 16");
    for (int i = 0; i < n_dst; ++i) DEXCT_REQUIRE(((uintptr_t)pp.p[i] & 15) == 0, "destination %d is not 16-byte aligned", i);
    const long long n16 = (long long)(bytes / 16);
    // Small CTAs on purpose: 128 threads x <= 32 registers fit into the register file beside three resident CTAs of the
    // decomposition kernel (3 x 256 x 80 registers), so the copy never keeps one of them from starting.
    const char *th = getenv("DEXCT_PUSH_THREADS");
    const int block = th && atoi(th) >= 32 && atoi(th) <= 512 ? atoi(th) / 32 * 32 : 128;
    long long grid = (n16 + block - 1) / block;
    const long long cap = max_ctas > 0 ? max_ctas : sm_count();
    if (grid > cap) grid = cap;
    peer_push_kernel<<<(unsigned)grid, block, 0, (cudaStream_t)stream>>>((const uint4 *)src, n16, pp, n_dst, skip);
    DEXCT_CUDA_OK(cudaGetLastError());
    return DEXCT_OK;
}

// The same transfer on the copy engines: one device-to-device DMA per destination, no SM time at all -- the choice when
// the copy runs beside a compute kernel that fills every SM (the decomposition's resident wave).
extern "C" int dexct_peer_push_dma(const void *src, size_t bytes, void *const *dst_ptrs_host, int32_t n_dst, int32_t skip,
                                   void *stream)
{
    if (int e = check_device()) return e;
    if (bytes == 0) return DEXCT_OK;
    DEXCT_REQUIRE(src && dst_ptrs_host && n_dst >= 1 && n_dst <= MAXP, "src and a table of 1..DEXCT_MAX_PEERS destinations required");
    for (int d = 0; d < n_dst; ++d) {
        if (d == skip) continue;
        DEXCT_REQUIRE(dst_ptrs_host[d], "destination %d is null", d);
        DEXCT_CUDA_OK(cudaMemcpyAsync(dst_ptrs_host[d], src, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    }
    return DEXCT_OK;
}
