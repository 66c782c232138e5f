// Version, error string and the one blocking read of libgsb.so.
#include <stdarg.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void gsb_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

GSB_API int gsb_version(void) { return 100; }

GSB_API const char* gsb_last_error(void) { return g_err; }

GSB_API int gsb_read_i64_host(const int64_t* src, int64_t* dst_host, void* stream) {
    GSB_CHECK_ARG(src && dst_host);
    GSB_CUDA(cudaMemcpyAsync(dst_host, src, sizeof(int64_t), cudaMemcpyDeviceToHost, gsb_stream(stream)));
    GSB_CUDA(cudaStreamSynchronize(gsb_stream(stream)));
    return 0;
}
