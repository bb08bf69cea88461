// common.cuh — error reporting shared by the translation units of libzebra_tppr.so.
#pragma once
#include <cuda_runtime.h>
#include <string>

namespace zebra {
int fail(const char *fmt, ...);          // records the message for zebra_last_error(), returns 1
}

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t _e = (call);                                                                   \
        if (_e != cudaSuccess)                                                                     \
            return zebra::fail("%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)
