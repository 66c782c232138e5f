// Per-unit timing trace of the compositing kernels (debug builds only: make TRACE=1).
// Every unit records (start ns, end ns, SM id, list length << 32 | kept entries); read back with
// the exported gsb_debug_trace_* functions by tools/trace_units.py.
#pragma once
#include "common.cuh"
#define GSB_TRACE_UNITS 65536
__device__ __forceinline__ unsigned long long gsb_gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ unsigned gsb_smid() {
    unsigned t;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(t));
    return t;
}
__device__ __forceinline__ void gsb_trace_unit(unsigned long long* buf, int unit, int lane,
                                               unsigned long long t_begin, int n, int kept) {
    if (lane == 0 && unit < GSB_TRACE_UNITS) {
        buf[unit * 4] = t_begin;
        buf[unit * 4 + 1] = gsb_gtimer();
        buf[unit * 4 + 2] = gsb_smid();
        buf[unit * 4 + 3] = ((unsigned long long)(unsigned)n << 32) | (unsigned)kept;
    }
}
#define GSB_TRACE_DEFINE(buf, fn)                                                        \
    __device__ unsigned long long buf[4 * GSB_TRACE_UNITS];                              \
    GSB_API int fn(unsigned long long* host, int n_units) {                              \
        cudaDeviceSynchronize();                                                         \
        return (int)cudaMemcpyFromSymbol(host, buf, sizeof(unsigned long long) * 4 * n_units); \
    }
