// Shared helpers of libdexct.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/dexct.h"

namespace dexct {

void set_error(const char *fmt, ...);
int check_device();          // DEXCT_OK or DEXCT_ENODEVICE (message set)
int sm_count();

// Exclusive use of one __constant__ table ("slot") of the library on the current device, across host threads and CUDA
// streams: the constructor takes a host mutex and makes `stream` wait for the previous user of the slot; the caller then
// enqueues  [fill staging] -> cudaMemcpyToSymbolAsync -> [kernel reading the table]  on `stream`; the destructor records
// the completion event and releases the mutex.  staging() is a library-owned 64 KiB device buffer of that slot.
enum { SLOT_GN = 0, SLOT_SIDDON = 1, SLOT_COUNT = 2 };
constexpr size_t SLOT_STAGING_BYTES = 64 * 1024;
class ConstSlot {
public:
    ConstSlot(int slot, cudaStream_t stream);
    ~ConstSlot();
    bool ok() const { return ok_; }
    void *staging() const { return staging_; }
private:
    int slot_, dev_;
    cudaStream_t stream_;
    bool ok_;
    void *staging_;
};

#define DEXCT_CUDA_OK(expr)                                                                        \
    do {                                                                                           \
        cudaError_t e__ = (expr);                                                                  \
        if (e__ != cudaSuccess) {                                                                  \
            dexct::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
            return DEXCT_ECUDA;                                                                    \
        }                                                                                          \
    } while (0)

#define DEXCT_REQUIRE(cond, ...)                                                                   \
    do {                                                                                           \
        if (!(cond)) {                                                                             \
            dexct::set_error(__VA_ARGS__);                                                         \
            return DEXCT_EINVAL;                                                                   \
        }                                                                                          \
    } while (0)

// Arithmetic with exactly one rounding per call and no compiler contraction.  The ray traversal is written with
// these so that its voxel-index sequence is reproducible operation-for-operation by oracle/siddon_oracle.c
// (section "B" there), in both precisions.
template <typename R> struct Rn;
template <> struct Rn<float> {
    static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
    static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
    static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
    static __device__ __forceinline__ float fma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
    static __device__ __forceinline__ float floor(float a) { return floorf(a); }
    static __device__ __forceinline__ float min(float a, float b) { return fminf(a, b); }
    static __device__ __forceinline__ float max(float a, float b) { return fmaxf(a, b); }
};
template <> struct Rn<double> {
    static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
    static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
    static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
    static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
    static __device__ __forceinline__ double fma(double a, double b, double c) { return __fma_rn(a, b, c); }
    static __device__ __forceinline__ double floor(double a) { return ::floor(a); }
    static __device__ __forceinline__ double min(double a, double b) { return fmin(a, b); }
    static __device__ __forceinline__ double max(double a, double b) { return fmax(a, b); }
};

}  // namespace dexct
