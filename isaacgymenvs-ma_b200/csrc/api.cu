// api.cu -- library-level entry points of the C ABI (include/ige_b200.h).
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>

#include "ige_common.cuh"

static thread_local char g_err[512] = "no error";
static std::atomic<uint64_t> g_launches{0};

int ige_fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

void ige_count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

int ige_check_launch(const char *what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return ige_fail((int)e, "%s: launch failed: %s", what, cudaGetErrorString(e));
    ige_count_launch();
    return IGE_OK;
}

extern "C" int ige_version(void) { return IGE_ABI_VERSION; }
extern "C" const char *ige_last_error_string(void) { return g_err; }
extern "C" uint64_t ige_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

// Debug knob: 0 = always take the generic (any-shape) kernels instead of the shape-specialised fast paths
// (ShadowHand 24/20, Franka MA pairs).  Also IGE_GENERIC=1.  Results are identical; tests run both.
static int g_fast_paths = -1;
extern "C" void ige_set_fast_paths(int on) { g_fast_paths = on ? 1 : 0; }
extern "C" int ige_fast_paths() {
    if (g_fast_paths < 0) {
        const char *e = getenv("IGE_GENERIC");
        g_fast_paths = (e && e[0] == '1') ? 0 : 1;
    }
    return g_fast_paths;
}

// Debug knob: 0 = plain stream-serialised launches instead of programmatic dependent launch (also IGE_PDL=0).
static int g_use_pdl = -1;
extern "C" void ige_set_use_pdl(int on) { g_use_pdl = on ? 1 : 0; }
extern "C" int ige_use_pdl() {
    if (g_use_pdl < 0) {
        const char *e = getenv("IGE_PDL");
        g_use_pdl = (e && e[0] == '0') ? 0 : 1;
    }
    return g_use_pdl;
}

// PCI bus id ("0000:1b:00.0") of a device: the host side binds each rank to its GPU's NUMA node before it allocates
// pinned arenas (distributed.bind_to_gpu_numa_node).
extern "C" int ige_device_pci_bus_id(int device, char *out, int len) {
    if (!out || len < 16) return ige_fail(IGE_ENULL, "ige_device_pci_bus_id: buffer too small");
    cudaError_t e = cudaDeviceGetPCIBusId(out, len, device);
    if (e != cudaSuccess) { cudaGetLastError(); return ige_fail((int)e, "cudaDeviceGetPCIBusId: %s", cudaGetErrorString(e)); }
    return IGE_OK;
}
