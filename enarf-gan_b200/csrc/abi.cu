// Error plumbing and argument checks of the C ABI (include/enarf_b200.h).
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void enarf_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* enarf_last_error(void) { return g_err; }
extern "C" int enarf_version(void) { return 153; }

// SM count of the CURRENT device (the one the caller's pointers live on), cached per device id: a process that
// drives several GPUs gets each one's own grid size.  Benign race: concurrent first calls store the same value.
int enarf_sm_count() {
    static int sms[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) {
        int v = 0;
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
        return v > 0 ? v : 148;
    }
    if (!sms[dev]) {
        int v = 0;
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
        sms[dev] = v > 0 ? v : 148;
    }
    return sms[dev];
}

int enarf_check_scene(const enarf_scene_t* sc, const char* who) {
    ENARF_REQUIRE(sc != nullptr, ENARF_ERR_INVALID_ARG, "%s: scene is NULL", who);
    ENARF_REQUIRE(sc->B > 0 && sc->P > 0 && sc->H > 1 && sc->W > 1, ENARF_ERR_INVALID_ARG,
                  "%s: bad scene sizes B=%d P=%d H=%d W=%d", who, sc->B, sc->P, sc->H, sc->W);
    ENARF_REQUIRE(sc->P <= ENARF_MAX_PARTS, ENARF_ERR_UNSUPPORTED, "%s: P=%d exceeds %d parts", who, sc->P,
                  ENARF_MAX_PARTS);
    ENARF_REQUIRE(sc->pose && sc->cscale && sc->canon_pose, ENARF_ERR_INVALID_ARG, "%s: NULL pose buffers", who);
    ENARF_REQUIRE(sc->feat_planes && sc->prob_planes, ENARF_ERR_INVALID_ARG, "%s: NULL plane buffers", who);
    ENARF_REQUIRE(sc->feat_dtype == ENARF_DTYPE_F32 || sc->feat_dtype == ENARF_DTYPE_BF16, ENARF_ERR_INVALID_ARG,
                  "%s: bad feat_dtype %d", who, sc->feat_dtype);
    ENARF_REQUIRE(((uintptr_t)sc->feat_planes & 15) == 0, ENARF_ERR_INVALID_ARG,
                  "%s: feat_planes must be 16-byte aligned", who);
    ENARF_REQUIRE(sc->w1 && sc->w2 && sc->w3 && sc->b1 && sc->b2 && sc->b3, ENARF_ERR_INVALID_ARG,
                  "%s: NULL decoder buffers", who);
    return ENARF_OK;
}

// Dynamic tile scheduling needs one zeroed int per launch.  The library must not allocate, so the counters are
// slots of a static device array (one copy per device: cudaGetSymbolAddress resolves it for the current device)
// handed out round-robin by a process-wide atomic cursor: launches stay independent unless more than
// ENARF_COUNTER_SLOTS of them are in flight at once.  This cursor and the SM-count cache above are the library's
// only mutable globals besides the thread-local error string.
#define ENARF_COUNTER_SLOTS 4096
__device__ int g_tile_counters[ENARF_COUNTER_SLOTS];
static unsigned g_next_slot = 0;

int* enarf_tile_counter(cudaStream_t stream) {
    int* base = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&base, g_tile_counters);
    if (e != cudaSuccess) {
        enarf_set_error("enarf_tile_counter: %s", cudaGetErrorString(e));
        return nullptr;
    }
    const unsigned slot = __atomic_fetch_add(&g_next_slot, 1u, __ATOMIC_RELAXED) % ENARF_COUNTER_SLOTS;
    e = cudaMemsetAsync(base + slot, 0, sizeof(int), stream);
    if (e != cudaSuccess) {
        enarf_set_error("enarf_tile_counter: %s", cudaGetErrorString(e));
        return nullptr;
    }
    return base + slot;
}
