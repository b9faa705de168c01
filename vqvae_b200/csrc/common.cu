// common.cu -- error text, device info, exclusive scan used by the CSR builders.
#include <stdarg.h>
#include "common.cuh"
#include "scan.cuh"

static thread_local char g_err[512] = "";

void geocb_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

GEOCB_API int geocb_abi_version(void) { return GEOCB_ABI_VERSION; }
GEOCB_API const char* geocb_last_error(void) { return g_err; }

int geocb_devinfo(DevInfo* out) {
    static thread_local int cached_dev = -1;
    static thread_local DevInfo cached;
    int dev = 0;
    CU_TRY(cudaGetDevice(&dev));
    if (dev != cached_dev) {
        int major = 0, minor = 0, sms = 0;
        CU_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
        CU_TRY(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
        CU_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        cached.sm_count = sms;
        cached.cc = major * 10 + minor;
        cached_dev = dev;
    }
    if (cached.cc / 10 != 10) {
        geocb_set_error("libgeocb is built for sm_100a only; device has compute capability %d", cached.cc);
        return GEOCB_ERR_UNSUPPORTED;
    }
    *out = cached;
    return 0;
}

GEOCB_API int geocb_device_info(int* sm_count, int* cc) {
    DevInfo di;
    int rc = geocb_devinfo(&di);
    if (rc) return rc;
    if (sm_count) *sm_count = di.sm_count;
    if (cc) *cc = di.cc;
    return 0;
}

// ------------------------------------------------------------------ scan ----
namespace {
constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ int64_t block_exclusive_scan(int64_t v, int64_t* total) {
    // v: this thread's sum; returns exclusive prefix over the block
    __shared__ int64_t warp_sums[kScanThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int64_t inc = v;
    for (int o = 1; o < 32; o <<= 1) {
        int64_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int64_t w = lane < kScanThreads / 32 ? warp_sums[lane] : 0;
        for (int o = 1; o < kScanThreads / 32; o <<= 1) {
            int64_t t = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += t;
        }
        if (lane < kScanThreads / 32) warp_sums[lane] = w;
    }
    __syncthreads();
    int64_t base = warp ? warp_sums[warp - 1] : 0;
    *total = warp_sums[kScanThreads / 32 - 1];
    __syncthreads();
    return base + inc - v;
}

template <typename T>
__global__ void scan_tile_sums(const T* in, int64_t n, int64_t* tile_sums) {
    const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
    int64_t s = 0;
    for (int i = 0; i < kScanItems; ++i)
        if (base + i < n) s += (int64_t)in[base + i];
    int64_t total;
    block_exclusive_scan(s, &total);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

__global__ void scan_of_tile_sums(int64_t* tile_sums, int64_t ntiles, int64_t* grand_total) {
    int64_t carry = 0;
    for (int64_t base = 0; base < ntiles; base += kScanThreads) {
        int64_t i = base + threadIdx.x;
        int64_t v = i < ntiles ? tile_sums[i] : 0, total;
        int64_t ex = block_exclusive_scan(v, &total);
        if (i < ntiles) tile_sums[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) *grand_total = carry;
}

template <typename T>
__global__ void scan_apply(const T* in, int64_t n, const int64_t* tile_sums, int64_t* out) {
    const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
    int64_t vals[kScanItems], s = 0;
    for (int i = 0; i < kScanItems; ++i) {
        vals[i] = base + i < n ? (int64_t)in[base + i] : 0;
        s += vals[i];
    }
    int64_t total;
    int64_t ex = block_exclusive_scan(s, &total) + tile_sums[blockIdx.x];
    for (int i = 0; i < kScanItems; ++i)
        if (base + i < n) {
            out[base + i] = ex;
            ex += vals[i];
        }
}

template <typename T>
int scan_impl(const T* in, int64_t n, int64_t* out, cudaStream_t st) {
    const int64_t ntiles = n > 0 ? (n + kScanTile - 1) / kScanTile : 0;
    if (n == 0) {
        CU_TRY(cudaMemsetAsync(out, 0, 8, st));
        return 0;
    }
    int64_t* tile_sums = nullptr;
    CU_TRY(cudaMallocAsync((void**)&tile_sums, (size_t)ntiles * 8, st));
    scan_tile_sums<T><<<(unsigned)ntiles, kScanThreads, 0, st>>>(in, n, tile_sums);
    scan_of_tile_sums<<<1, kScanThreads, 0, st>>>(tile_sums, ntiles, out + n);
    scan_apply<T><<<(unsigned)ntiles, kScanThreads, 0, st>>>(in, n, tile_sums, out);
    KERNEL_CHECK();
    CU_TRY(cudaFreeAsync(tile_sums, st));
    return 0;
}
}  // namespace

int scan_exclusive_i32(const int32_t* in, int64_t n, int64_t* out, cudaStream_t st) { return scan_impl(in, n, out, st); }
int scan_exclusive_u8(const uint8_t* in, int64_t n, int64_t* out, cudaStream_t st) { return scan_impl(in, n, out, st); }
