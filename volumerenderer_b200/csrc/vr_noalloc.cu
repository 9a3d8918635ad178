// vr_noalloc.cu -- vr_light_gen_static: the light-generation pass for callers that must not allocate device memory.
//
// Who needs this. vr_dispatch_light_gen keeps its working memory (one slab of candidate lights per photon, the per-photon
// counts and their scan) in buffers the context grows with cudaMalloc. Two kinds of host cannot live with that:
//   * a host that handed the whole GPU to its own memory pool (a DCC application's arena, a framework's caching allocator
//     with a hard memory limit): a cudaMalloc issued from inside a library either fails there or escapes the host's
//     accounting, and such hosts usually forbid it for plug-ins outright;
//   * a host that calls from the middle of a frame in flight: cudaMalloc and cudaFree may wait for the whole device, which
//     stalls every other stream of the process (video encode, UI, another renderer) for the duration.
// For them this entry point runs the same kernels with the working memory in ONE fixed array that is part of the library
// image (g_lg_scratch below, about 2.1 MB, present on every GPU that loads libvr until the process exits). The caller
// brings the volume, the light list and the counter as device pointers out of its own pool; nothing is allocated here, by
// this file or by anything it calls.
//
// What the fixed array costs: a hard limit of VR_STATIC_MAX_PHOTONS photons per call, and no concurrency -- every call on a
// GPU uses the same array, so two calls whose kernels can overlap on that GPU (different streams, or different host
// threads) corrupt each other's lights. Calls on ONE stream are safe back to back.
//
// The array's name is only a handle on the host side; kernels need its device address, which differs per GPU. The first
// call on each GPU asks the runtime for it (cudaGetSymbolAddress with that GPU current) and keeps it; this also makes the
// runtime load this file's device code on that GPU. Make that first call during start-up, outside any stream capture.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <mutex>

#include "vr_kernels.cuh"

#define VR_STATIC_MAX_PHOTONS_ 4096u
static_assert(VR_STATIC_MAX_PHOTONS_ == VR_STATIC_MAX_PHOTONS, "include/vr.h and vr_noalloc.cu disagree on the photon limit");

namespace {

// Layout of the fixed working memory (what vrk_lightgen calls counts / offsets / total / scan_tmp / slab). The slab holds
// 16 candidate lights of two float4 per photon: vrk_lightgen_local_capacity(), checked at set-up.
constexpr uint32_t kSlabLightsPerPhoton = 16;
struct LgScratch {
    unsigned long long total;                                    // scan grand total: cleared at the start of every call
    unsigned long long scan_tmp[VR_STATIC_MAX_PHOTONS_ / 4096 + 2];
    unsigned long long offsets[VR_STATIC_MAX_PHOTONS_];
    uint32_t counts[VR_STATIC_MAX_PHOTONS_];
    float4 slab[(size_t)VR_STATIC_MAX_PHOTONS_ * kSlabLightsPerPhoton * 2];
};

}  // namespace

// The working memory itself. In host code `g_lg_scratch` is NOT a device address (see lg_scratch_on below).
static __device__ __align__(256) unsigned char g_lg_scratch[sizeof(LgScratch)];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;
LgScratch* g_scratch_addr[kMaxDevices];  // device address of g_lg_scratch per GPU, nullptr until that GPU's first call

int fail(char* text, uint64_t text_size, int code, const char* fmt, ...) {
    if (text && text_size) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(text, (size_t)text_size, fmt, ap);
        va_end(ap);
    }
    return code;
}

// Device address of the working memory on `device`, which must be the current device. Looked up once per GPU.
cudaError_t lg_scratch_on(int device, LgScratch** out) {
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_scratch_addr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_lg_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_addr[device] = static_cast<LgScratch*>(p);
    }
    *out = g_scratch_addr[device];
    return cudaSuccess;
}

template <class T>
T field(const uint8_t* b, uint64_t off) {
    T v;
    memcpy(&v, b + off, sizeof(T));
    return v;
}

// The facts the kernels take by value, read from a host copy of the grid's first bytes (GridData 672 B, TreeData 64 B,
// then wherever the tree says the root is): the same fields vr_set_grid_device reads.
int describe_grid(const void* device_bytes, uint64_t size, const uint8_t* h, uint64_t have, VrGridDev* out, char* text,
                  uint64_t text_size) {
    if (size < 800 || size % 32 != 0)
        return fail(text, text_size, VR_E_BADGRID, "vr_light_gen_static: %llu bytes cannot be a NanoVDB grid", (unsigned long long)size);
    if (have > size) have = size;
    if (have < 736) return fail(text, text_size, VR_E_ARG, "vr_light_gen_static: grid header too short (%llu bytes)", (unsigned long long)have);
    const uint64_t magic = field<uint64_t>(h, 0);
    const bool known = magic == 0x304244566f6e614eULL || magic == 0x314244566f6e614eULL || magic == 0x324244566f6e614eULL;
    if (!known || field<uint32_t>(h, 636) != 1u)
        return fail(text, text_size, VR_E_BADGRID, "vr_light_gen_static: not a NanoVDB float grid");
    if (field<uint64_t>(h, 32) > size) return fail(text, text_size, VR_E_BADGRID, "vr_light_gen_static: gridSize field exceeds the buffer");
    const uint64_t root = 672 + field<uint64_t>(h, 672 + 24);
    if (root + 64 > size) return fail(text, text_size, VR_E_BADGRID, "vr_light_gen_static: root node outside the buffer");
    if (root + 64 > have)
        return fail(text, text_size, VR_E_ARG, "vr_light_gen_static: grid header too short, the root node ends at byte %llu",
                    (unsigned long long)(root + 64));
    const uint32_t tiles = field<uint32_t>(h, root + 24);
    if (root + 64 + 32ull * tiles > size) return fail(text, text_size, VR_E_BADGRID, "vr_light_gen_static: root tile table outside the buffer");
    VrGridDev g{};
    g.base = static_cast<const uint8_t*>(device_bytes);
    g.root = root;
    g.tile_count = tiles;
    for (int a = 0; a < 3; a++) {
        g.bmin[a] = field<int32_t>(h, root + 4 * a);
        g.bmax[a] = field<int32_t>(h, root + 12 + 4 * a);
        g.vec[a] = field<float>(h, 296 + 72 + 4 * a);
    }
    for (int a = 0; a < 9; a++) {
        g.mat[a] = field<float>(h, 296 + 4 * a);
        g.inv[a] = field<float>(h, 296 + 36 + 4 * a);
    }
    *out = g;
    return VR_OK;
}

// The photon walk advances t by the step in FP32 and restarts at 0 on a scatter, so (farthest box corner as seen from the
// light) / step bounds its length; beyond ~8e6 steps t + step == t and the loop would never end.
bool walk_terminates(const VrGridDev& g, const vr_ubo& u) {
    if (!(u.rayMarchingStepSize > 0.0f)) return false;
    double far2 = 0.0;
    const double s[3] = {(double)u.lightSourceWorldPos[0] - g.vec[0], (double)u.lightSourceWorldPos[1] - g.vec[1],
                         (double)u.lightSourceWorldPos[2] - g.vec[2]};
    for (int a = 0; a < 3; a++) {
        const double ci = s[0] * g.inv[3 * a] + s[1] * g.inv[3 * a + 1] + s[2] * g.inv[3 * a + 2];
        const double d = fmax(fabs(ci - (double)g.bmin[a]), fabs((double)g.bmax[a] + 1.0 - ci));
        far2 += d * d;
    }
    double reach = sqrt(far2);
    const double maxd = (double)u.rayMaxDistance;
    if (maxd == maxd && maxd < reach) reach = maxd > 0.0 ? maxd : 0.0;
    return reach / (double)u.rayMarchingStepSize <= 8.0e6;
}

}  // namespace

extern "C" int vr_light_gen_static(int cuda_device, void* cuda_stream, const void* grid_device_bytes, uint64_t grid_size,
                                   const void* grid_header, uint64_t grid_header_size, const vr_ubo* params, uint32_t gx,
                                   uint32_t gy, uint32_t gz, vr_lightgen_variant variant, vr_ray_light* lights_device,
                                   void* counter_device, char* error_text, uint64_t error_text_size) {
    char* et = error_text;
    const uint64_t es = error_text_size;
    if (et && es) et[0] = 0;
    if (cuda_device < 0 || cuda_device >= kMaxDevices) return fail(et, es, VR_E_ARG, "vr_light_gen_static: device %d out of range", cuda_device);
    if (!grid_device_bytes || !grid_header || !params || !lights_device || !counter_device)
        return fail(et, es, VR_E_ARG, "vr_light_gen_static: NULL argument");
    if (((uintptr_t)grid_device_bytes & 31) || ((uintptr_t)lights_device & 15) || ((uintptr_t)counter_device & 7))
        return fail(et, es, VR_E_ARG, "vr_light_gen_static: the grid must be 32-byte, the light list 16-byte and the counter 8-byte aligned");
    if (gx == 0 || gy == 0 || gz == 0) return fail(et, es, VR_E_ARG, "vr_light_gen_static: dispatch dimensions must be >= 1");
    const uint64_t nphot = (uint64_t)gx * 4 * (uint64_t)gy * 4 * (uint64_t)gz;  // work-groups of 4x4x1 photons (light_gen.comp:7)
    if (gx > VR_STATIC_MAX_PHOTONS_ || gy > VR_STATIC_MAX_PHOTONS_ || gz > VR_STATIC_MAX_PHOTONS_ || nphot > VR_STATIC_MAX_PHOTONS_)
        return fail(et, es, VR_E_CAPACITY,
                    "vr_light_gen_static: %llu x %llu x %llu work-groups are more than %u photons, the most the fixed working memory "
                    "holds; use vr_dispatch_light_gen (which allocates) for larger dispatches",
                    (unsigned long long)gx, (unsigned long long)gy, (unsigned long long)gz, VR_STATIC_MAX_PHOTONS_);
    if (variant != VR_LG_RESET_ORIGIN && variant != VR_LG_KEEP_ORIGIN)
        return fail(et, es, VR_E_ARG, "vr_light_gen_static: unknown light-generation variant %d", (int)variant);
    VrGridDev g{};
    const int rc = describe_grid(grid_device_bytes, grid_size, static_cast<const uint8_t*>(grid_header), grid_header_size, &g, et, es);
    if (rc != VR_OK) return rc;
    if (!walk_terminates(g, *params))
        return fail(et, es, VR_E_ARG, "vr_light_gen_static: rayMarchingStepSize %g gives a photon walk that does not end on this grid",
                    (double)params->rayMarchingStepSize);

    cudaError_t e = cudaSetDevice(cuda_device);
    if (e != cudaSuccess) return fail(et, es, VR_E_CUDA, "vr_light_gen_static: cudaSetDevice(%d): %s", cuda_device, cudaGetErrorString(e));
    LgScratch* s = nullptr;
    e = lg_scratch_on(cuda_device, &s);
    if (e != cudaSuccess)
        return fail(et, es, VR_E_CUDA, "vr_light_gen_static: cudaGetSymbolAddress on device %d: %s", cuda_device, cudaGetErrorString(e));
    if (vrk_lightgen_local_capacity() != kSlabLightsPerPhoton)
        return fail(et, es, VR_E_STATE, "vr_light_gen_static: internal: slab sized for %u lights per photon, the kernels use %u",
                    kSlabLightsPerPhoton, vrk_lightgen_local_capacity());

    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    // The array still holds the previous call's data. The kernels write counts, offsets, scan_tmp and every slab entry they
    // later read before reading them; the grand total is cleared here so that no path can pick up a stale value.
    e = cudaMemsetAsync(&s->total, 0, sizeof(s->total), st);
    // what vr_reset_lights does for a context (vkCmdFillBuffer of the counter, src/main.cpp:722-728): the list starts empty
    if (e == cudaSuccess) e = cudaMemsetAsync(counter_device, 0, sizeof(VrLightCounter), st);
    if (e != cudaSuccess) return fail(et, es, VR_E_CUDA, "vr_light_gen_static: cudaMemsetAsync: %s", cudaGetErrorString(e));
    vrk_lightgen(st, g, *params, variant == VR_LG_KEEP_ORIGIN ? 1 : 0, gx * 4, gy * 4, gz, s->counts, s->offsets, &s->total, s->scan_tmp,
                 s->slab, lights_device, static_cast<VrLightCounter*>(counter_device));
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(et, es, VR_E_CUDA, "vr_light_gen_static: kernel launch: %s", cudaGetErrorString(e));
    return VR_OK;
}
