// Library-level entry points of libqdc_b200.so: version, error string, device info.
#include <stdarg.h>
#include <string.h>
#include "qdc_common.cuh"

static thread_local char g_err[1024] = "";

void qdc_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int qdc_sm_count() {
    static thread_local int cached_dev = -1, cached = QDC_NUM_SMS_B200;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return QDC_NUM_SMS_B200;
    if (dev != cached_dev) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
            cached = n;
        cached_dev = dev;
    }
    return cached;
}

extern "C" int qdc_version(void) { return 100; }

extern "C" const char* qdc_last_error(void) { return g_err; }

extern "C" int qdc_device_info(int* sm_count, int* cc_major, int* cc_minor) {
    int dev = 0;
    QDC_CHECK_CUDA(cudaGetDevice(&dev));
    if (sm_count) QDC_CHECK_CUDA(cudaDeviceGetAttribute(sm_count, cudaDevAttrMultiProcessorCount, dev));
    if (cc_major) QDC_CHECK_CUDA(cudaDeviceGetAttribute(cc_major, cudaDevAttrComputeCapabilityMajor, dev));
    if (cc_minor) QDC_CHECK_CUDA(cudaDeviceGetAttribute(cc_minor, cudaDevAttrComputeCapabilityMinor, dev));
    return QDC_OK;
}
