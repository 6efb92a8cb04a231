// core.cu -- library plumbing of libolbank.so: errors, device checks, pinned memory, replayable noise,
// TMA descriptor encoding.  No DSP here.
#include <cuda.h>
#include <sched.h>
#include <stdarg.h>

#include "common.cuh"
#include "tmap.h"

namespace olb {

static thread_local char t_err[512] = "";
std::atomic<uint64_t> g_launches{0};

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof(t_err), fmt, ap);
    va_end(ap);
}

int fail_cuda(cudaError_t e, const char *what, const char *file, int line) {
    set_error("CUDA error %d (%s) in %s at %s:%d", (int) e, cudaGetErrorString(e), what, file, line);
    return OLB_E_CUDA;
}

// The library ships sm_100a SASS only; refuse anything else up front instead of failing at launch.
int require_sm100(int device) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        set_error("no CUDA device available (%s); libolbank has no CPU fallback", cudaGetErrorString(e));
        return OLB_E_CUDA;
    }
    OLB_ARG(device >= 0 && device < n, "device %d out of range (have %d)", device, n);
    int major = 0;
    OLB_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    if (major != 10) {
        set_error("device %d is sm_%dx; libolbank is built for sm_100a only", device, major);
        return OLB_E_CUDA;
    }
    return OLB_OK;
}

// ---- TMA descriptor for a frame-interleaved block [frames][units] fp32 --------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn) p;
    }
    return fn;
}

int make_tmap_frames(TensorMap *out, const float *base, uint64_t frames, uint64_t units, uint32_t box_units, uint32_t box_frames) {
    static_assert(sizeof(TensorMap) == sizeof(CUtensorMap), "TensorMap must mirror CUtensorMap");
    EncodeTiledFn enc = get_encode();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled not available from the driver");
        return OLB_E_CUDA;
    }
    OLB_ARG(((uintptr_t) base & 15) == 0, "audio buffer must be 16-byte aligned for the TMA path");
    OLB_ARG((units & 3) == 0, "units must be a multiple of 4 for the TMA path");
    cuuint64_t dims[2] = {units, frames};
    cuuint64_t strides[1] = {units * sizeof(float)};
    cuuint32_t box[2] = {box_units, box_frames};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc((CUtensorMap *) out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *) base, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult %d (units=%llu frames=%llu box=%ux%u)", (int) r,
                  (unsigned long long) units, (unsigned long long) frames, box_units, box_frames);
        return OLB_E_CUDA;
    }
    return OLB_OK;
}

// ---- replayable noise ---------------------------------------------------------------------------
// One thread per (unit, 4-frame Philox block).  FRAMES layout: adjacent threads = adjacent units, so
// each of the four stores is a coalesced row segment.  PLANAR: adjacent threads = adjacent 4-frame
// blocks of one unit (16 B per thread, contiguous).
__global__ void k_noise_frames(float *out, uint32_t seed, uint32_t first_unit, uint32_t units, uint64_t first_frame,
                               uint64_t frames, float scale, uint64_t ld) {
    const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= units) return;
    const uint64_t b0 = first_frame >> 2, b1 = (first_frame + frames + 3) >> 2;
    for (uint64_t b = b0 + blockIdx.y; b < b1; b += gridDim.y) {
        uint32_t w[4];
        philox4x32_10(seed, first_unit + u, b, w);
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint64_t f = b * 4 + j;
            if (f >= first_frame && f < first_frame + frames) out[(f - first_frame) * ld + u] = noise_from_u32(w[j]) * scale;
        }
    }
}

__global__ void k_noise_planar(float *out, uint32_t seed, uint32_t first_unit, uint32_t units, uint64_t first_frame,
                               uint64_t frames, float scale) {
    const uint64_t b0 = first_frame >> 2, nb = ((first_frame + frames + 3) >> 2) - b0;
    const uint32_t u = blockIdx.y;
    for (uint64_t i = blockIdx.x * (uint64_t) blockDim.x + threadIdx.x; i < nb; i += (uint64_t) gridDim.x * blockDim.x) {
        uint32_t w[4];
        philox4x32_10(seed, first_unit + u, b0 + i, w);
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint64_t f = (b0 + i) * 4 + j;
            if (f >= first_frame && f < first_frame + frames) out[(uint64_t) u * frames + (f - first_frame)] = noise_from_u32(w[j]) * scale;
        }
    }
}

}  // namespace olb

using namespace olb;

OLB_API const char *olb_last_error(void) { return t_err; }
OLB_API const char *olb_version(void) { return "olbank 0.1 (sm_100a)"; }
OLB_API uint64_t olb_launch_count(void) { return g_launches.load(); }

OLB_API int olb_device_count(int *count) {
    OLB_ARG(count, "count is NULL");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        cudaGetLastError();
        n = 0;
    }
    *count = n;
    return OLB_OK;
}

// ---- host placement: the NUMA node a GPU hangs off (sysfs), the calling thread bound to that node's cores ----------
// Host-buffer entry points move every block over PCIe in both directions; with one process per GPU on a two-socket box the
// copies of a rank whose buffers live on the other socket cross the inter-socket link and all ranks queue on one
// memory controller.  olb_bind_thread_near pins the CALLING thread to the cores of the GPU's node (so that the pages it
// allocates and touches afterwards land there); olb_host_alloc_near = bind + pinned allocation.
static int gpu_numa_node(int device, int *node) {
    char bdf[32] = "";
    OLB_CUDA(cudaDeviceGetPCIBusId(bdf, sizeof(bdf), device));
    for (char *p = bdf; *p; p++)
        if (*p >= 'A' && *p <= 'F') *p = (char) (*p - 'A' + 'a');
    char path[128];
    snprintf(path, sizeof(path), "/sys/bus/pci/devices/%s/numa_node", bdf);
    *node = -1;
    if (FILE *f = fopen(path, "r")) {
        if (fscanf(f, "%d", node) != 1) *node = -1;
        fclose(f);
    }
    return OLB_OK;
}

OLB_API int olb_bind_thread_near(int device, int *numa_node_out) {
    int node = -1;
    int rc = gpu_numa_node(device, &node);
    if (rc) return rc;
    if (numa_node_out) *numa_node_out = node;
    if (node < 0) return OLB_OK;  // the platform reports no locality: leave the thread where it is
    char path[128];
    snprintf(path, sizeof(path), "/sys/devices/system/node/node%d/cpulist", node);
    FILE *f = fopen(path, "r");
    if (!f) return OLB_OK;
    cpu_set_t set;
    CPU_ZERO(&set);
    int a = 0, b = 0, n = 0;
    char sep = 0;
    while (fscanf(f, "%d", &a) == 1) {  // "0-31,64-95"
        b = a;
        if (fscanf(f, "%c", &sep) == 1 && sep == '-') {
            if (fscanf(f, "%d", &b) != 1) b = a;
            if (fscanf(f, "%c", &sep) != 1) sep = 0;
        }
        for (int c = a; c <= b && c < CPU_SETSIZE; c++, n++) CPU_SET(c, &set);
        if (sep != ',') break;
    }
    fclose(f);
    if (n) {
        cpu_set_t allowed;  // stay inside what the container / launcher allows
        if (sched_getaffinity(0, sizeof(allowed), &allowed) == 0) {
            cpu_set_t both;
            CPU_AND(&both, &set, &allowed);
            if (CPU_COUNT(&both) > 0) sched_setaffinity(0, sizeof(both), &both);
        }
    }
    return OLB_OK;
}

OLB_API int olb_host_alloc_near(int device, void **ptr, size_t bytes, int *numa_node_out) {
    OLB_ARG(ptr && bytes, "olb_host_alloc_near: bad argument");
    int rc = olb_bind_thread_near(device, numa_node_out);
    if (rc) return rc;
    DeviceGuard g(device);
    OLB_CUDA(cudaHostAlloc(ptr, bytes, cudaHostAllocDefault));
    return OLB_OK;
}

OLB_API int olb_host_alloc(void **ptr, size_t bytes) {
    OLB_ARG(ptr && bytes, "olb_host_alloc: bad argument");
    OLB_CUDA(cudaHostAlloc(ptr, bytes, cudaHostAllocDefault));
    return OLB_OK;
}
OLB_API int olb_host_free(void *ptr) {
    if (ptr) OLB_CUDA(cudaFreeHost(ptr));
    return OLB_OK;
}

OLB_API int olb_noise_fill(float *out_dev, uint32_t seed, uint32_t first_unit, uint32_t units, uint64_t first_frame,
                           uint64_t frames, float scale, int layout, void *stream) {
    OLB_ARG(out_dev, "olb_noise_fill: out_dev is NULL");
    OLB_ARG(layout == OLB_LAYOUT_FRAMES || layout == OLB_LAYOUT_PLANAR, "olb_noise_fill: bad layout %d", layout);
    if (units == 0 || frames == 0) return OLB_OK;
    cudaStream_t st = (cudaStream_t) stream;
    const uint64_t nb = ((first_frame + frames + 3) >> 2) - (first_frame >> 2);
    if (layout == OLB_LAYOUT_FRAMES) {
        dim3 grid((units + 127) / 128, (unsigned) (nb < 1024 ? nb : 1024));
        k_noise_frames<<<grid, 128, 0, st>>>(out_dev, seed, first_unit, units, first_frame, frames, scale, units);
    } else {
        OLB_ARG(units <= 65535, "olb_noise_fill: planar layout supports at most 65535 units per call");
        uint64_t gx = (nb + 255) / 256;
        dim3 grid((unsigned) (gx < 4096 ? gx : 4096), units);
        k_noise_planar<<<grid, 256, 0, st>>>(out_dev, seed, first_unit, units, first_frame, frames, scale);
    }
    OLB_LAUNCHED();
    return OLB_OK;
}
