// vr_api.cu -- the C ABI of include/vr.h: a headless context that plays the role of the descriptor set + command
// buffer of VolumeApp::drawFrame (src/main.cpp:712-837) on one CUDA stream. No exceptions cross the ABI.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <new>
#include <string>

#include "../../include/vr.h"
#include "vr_kernels.cuh"

static_assert(sizeof(vr_ubo) == 96, "UBO must match the std140 block of shaders/common_bindings.h:19-34");
static_assert(sizeof(vr_ray_light) == 32, "RayLight std430 stride (shaders/common_bindings.h:10-12)");

namespace {
thread_local std::string g_create_error;

template <class T>
struct DevBuf {
    T* p = nullptr;
    uint64_t cap = 0;  // elements
    cudaError_t reserve(uint64_t n, bool keep = false, cudaStream_t st = 0) {
        if (n <= cap) return cudaSuccess;
        uint64_t ncap = n + n / 8 + 64;
        T* q = nullptr;
        cudaError_t e = cudaMalloc(&q, ncap * sizeof(T));
        if (e != cudaSuccess) return e;
        if (p) {
            if (keep) cudaMemcpyAsync(q, p, cap * sizeof(T), cudaMemcpyDeviceToDevice, st);
            cudaStreamSynchronize(st);
            cudaFree(p);
        }
        p = q;
        cap = ncap;
        return cudaSuccess;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

struct MarchKey {
    uint64_t grid_gen = 0;
    vr_tile tile = {0, 0, 0, 0};
    uint32_t dim[2] = {0, 0};
    float cam[3] = {0, 0, 0}, fov = 0, maxd = 0, step = 0, absorb = 0;
    bool valid = false;
    bool operator==(const MarchKey& o) const {
        return valid && o.valid && grid_gen == o.grid_gen && memcmp(&tile, &o.tile, sizeof(tile)) == 0 &&
               memcmp(dim, o.dim, sizeof(dim)) == 0 && memcmp(cam, o.cam, sizeof(cam)) == 0 && fov == o.fov &&
               maxd == o.maxd && step == o.step && absorb == o.absorb;
    }
};
}  // namespace

struct vr_ctx {
    int device = 0;
    cudaStream_t stream = nullptr, own_stream = nullptr;
    unsigned long long launch_base = 0;
    std::string err;
    // binding 2
    uint8_t* grid = nullptr;
    uint64_t grid_size = 0, grid_gen = 0, grid_cap = 0;
    VrGridDev gdev{};
    bool have_grid = false;
    // binding 5
    vr_ubo ubo{};
    // bindings 3/4
    DevBuf<vr_ray_light> lights;
    uint32_t light_capacity = 0;
    VrLightCounter* counter = nullptr;
    DevBuf<uint32_t> photon_counts;
    DevBuf<unsigned long long> photon_offsets;
    unsigned long long* photon_total = nullptr;
    uint32_t lg_groups[3] = {1, 1, 1};
    vr_lightgen_variant lg_variant = VR_LG_RESET_ORIGIN;
    // binding 0 (+ float accumulator and last frame colour)
    uint32_t width = 0, height = 0;
    float4* accum = nullptr;
    uchar4* rgba8 = nullptr;
    float* frame_color = nullptr;
    // gather work buffers
    DevBuf<uint32_t> px_counts;
    DevBuf<unsigned long long> px_offsets;
    DevBuf<unsigned long long> scan_tmp;
    DevBuf<VrSample> samples;
    DevBuf<float> contrib;
    DevBuf<VrEvalLight> eval_lights;
    DevBuf<float4> eval_lights2;
    int eval_variant = 2;  // 0: scalar kernel, 1: packed FP32x2 kernel, 2/3: packed speculative kernel (point/ray lists)
    DevBuf<uint32_t> light_steps;
    DevBuf<unsigned long long> light_offsets;
    unsigned long long* totals = nullptr;       // [0] samples, [1] expanded lights (device)
    unsigned long long* totals_host = nullptr;  // pinned
    VrWorkCounters* wc = nullptr;
    VrWorkCounters* wc_host = nullptr;  // pinned
    unsigned long long* eval_redo = nullptr;  // CTAs of the speculative eval kernel that repeated with the guard (cumulative)
    vr_counters last_counters{};
    bool march_cache = false;
    MarchKey march_key;
    uint64_t cached_samples = 0;
    // timing
    cudaEvent_t ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t tev[2] = {nullptr, nullptr};
    vr_timings timings{};
    bool timings_pending = false, lg_timed = false;
};

namespace {

int fail(vr_ctx* c, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (c) c->err = buf;
    else g_create_error = buf;
    return code;
}
#define VR_CUDA(c, call)                                                                                  \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess) return fail((c), VR_E_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

template <class T>
T rd(const uint8_t* b, uint64_t off) {
    T v;
    memcpy(&v, b + off, sizeof(T));
    return v;
}

int check_stream(vr_ctx* c) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(c, VR_E_CUDA, "kernel launch failed: %s", cudaGetErrorString(e));
    return VR_OK;
}

vr_tile full_tile(const vr_ctx* c) { return vr_tile{0, 0, c->width, c->height}; }

int validate_tile(vr_ctx* c, const vr_tile* t, vr_tile* out) {
    if (!t) {
        *out = full_tile(c);
        return VR_OK;
    }
    if (t->w == 0 || t->h == 0 || (uint64_t)t->x0 + t->w > c->width || (uint64_t)t->y0 + t->h > c->height)
        return fail(c, VR_E_ARG, "tile (%u,%u %ux%u) outside the %ux%u image", t->x0, t->y0, t->w, t->h, c->width, c->height);
    *out = *t;
    return VR_OK;
}

int finalize_timings(vr_ctx* c) {
    if (!c->timings_pending) return VR_OK;
    VR_CUDA(c, cudaEventSynchronize(c->ev[7]));
    float ms = 0;
    if (c->lg_timed) {
        cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]);
        c->timings.light_gen_ms = ms;
    }
    cudaEventElapsedTime(&ms, c->ev[2], c->ev[3]);
    c->timings.prepare_ms = ms;
    cudaEventElapsedTime(&ms, c->ev[3], c->ev[4]);
    float m1 = ms;
    cudaEventElapsedTime(&ms, c->ev[4], c->ev[5]);
    c->timings.march_ms = m1 + ms;
    cudaEventElapsedTime(&ms, c->ev[5], c->ev[6]);
    c->timings.eval_ms = ms;
    cudaEventElapsedTime(&ms, c->ev[6], c->ev[7]);
    c->timings.resolve_ms = ms;
    cudaEventElapsedTime(&ms, c->lg_timed ? c->ev[0] : c->ev[2], c->ev[7]);
    c->timings.total_ms = ms;
    c->timings_pending = false;
    return VR_OK;
}

}  // namespace

extern "C" {

const char* vr_version(void) { return "libvr 0.1 (sm_100a)"; }

const char* vr_last_error(const vr_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int vr_create(int cuda_device, vr_ctx** out) {
    if (!out) return fail(nullptr, VR_E_ARG, "vr_create: out is NULL");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(nullptr, VR_E_CUDA, "no CUDA device available (%s): libvr has no CPU fallback", cudaGetErrorString(e));
    if (cuda_device < 0 || cuda_device >= ndev) return fail(nullptr, VR_E_ARG, "cuda_device %d out of range [0,%d)", cuda_device, ndev);
    vr_ctx* c = new (std::nothrow) vr_ctx();
    if (!c) return fail(nullptr, VR_E_CUDA, "out of host memory");
    c->device = cuda_device;
#define VR_CREATE_CUDA(call)                                                                    \
    do {                                                                                        \
        cudaError_t e_ = (call);                                                                \
        if (e_ != cudaSuccess) {                                                                \
            fail(nullptr, VR_E_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_));            \
            vr_destroy(c);                                                                      \
            return VR_E_CUDA;                                                                   \
        }                                                                                       \
    } while (0)
    VR_CREATE_CUDA(cudaSetDevice(cuda_device));
    VR_CREATE_CUDA(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
    c->stream = c->own_stream;
    c->launch_base = vrk_launch_count();
    VR_CREATE_CUDA(cudaMalloc(&c->counter, sizeof(VrLightCounter)));
    VR_CREATE_CUDA(cudaMemsetAsync(c->counter, 0, sizeof(VrLightCounter), c->stream));
    VR_CREATE_CUDA(cudaMalloc(&c->photon_total, sizeof(unsigned long long)));
    VR_CREATE_CUDA(cudaMalloc(&c->totals, 2 * sizeof(unsigned long long)));
    VR_CREATE_CUDA(cudaMallocHost(&c->totals_host, 4 * sizeof(unsigned long long)));
    VR_CREATE_CUDA(cudaMalloc(&c->wc, sizeof(VrWorkCounters)));
    VR_CREATE_CUDA(cudaMallocHost(&c->wc_host, sizeof(VrWorkCounters)));
    VR_CREATE_CUDA(cudaMalloc(&c->eval_redo, sizeof(unsigned long long)));
    VR_CREATE_CUDA(cudaMemsetAsync(c->eval_redo, 0, sizeof(unsigned long long), c->stream));
    for (int i = 0; i < 8; i++) VR_CREATE_CUDA(cudaEventCreate(&c->ev[i]));
    for (int i = 0; i < 2; i++) VR_CREATE_CUDA(cudaEventCreate(&c->tev[i]));
#undef VR_CREATE_CUDA
    vr_default_params(&c->ubo, 0, 0);
    // the reference allocates 100000 * sizeof(RayLight host = 28 B) = 87500 lights at the 32 B GPU stride
    // (src/main.cpp:383-389)
    int rc = vr_set_max_lights_capacity(c, 87500);
    if (rc != VR_OK) {
        g_create_error = c->err;
        vr_destroy(c);
        return rc;
    }
    *out = c;
    return VR_OK;
}

int vr_destroy(vr_ctx* c) {
    if (!c) return VR_OK;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->grid) cudaFree(c->grid);
    c->lights.release();
    c->photon_counts.release();
    c->photon_offsets.release();
    c->px_counts.release();
    c->px_offsets.release();
    c->scan_tmp.release();
    c->samples.release();
    c->contrib.release();
    c->eval_lights.release();
    c->eval_lights2.release();
    c->light_steps.release();
    c->light_offsets.release();
    if (c->counter) cudaFree(c->counter);
    if (c->photon_total) cudaFree(c->photon_total);
    if (c->totals) cudaFree(c->totals);
    if (c->totals_host) cudaFreeHost(c->totals_host);
    if (c->wc) cudaFree(c->wc);
    if (c->wc_host) cudaFreeHost(c->wc_host);
    if (c->eval_redo) cudaFree(c->eval_redo);
    if (c->accum) cudaFree(c->accum);
    if (c->rgba8) cudaFree(c->rgba8);
    if (c->frame_color) cudaFree(c->frame_color);
    for (int i = 0; i < 8; i++)
        if (c->ev[i]) cudaEventDestroy(c->ev[i]);
    for (int i = 0; i < 2; i++)
        if (c->tev[i]) cudaEventDestroy(c->tev[i]);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
    return VR_OK;
}

void vr_default_params(vr_ubo* p, uint32_t width, uint32_t height) {
    if (!p) return;
    memset(p, 0, sizeof(*p));
    // src/main.cpp:546-559
    p->frameCount = 0;
    p->framebufferDim[0] = width, p->framebufferDim[1] = height;
    p->cameraPos[0] = 0.0f, p->cameraPos[1] = 20.0f, p->cameraPos[2] = -75.0f;
    p->fov = 45.0f;
    p->photonInitialIntensity = 100.0f;
    p->scatteringProbability = 0.05f;
    p->absorptionCoefficient = 0.05f;
    p->maxLights = 1000;
    p->rayMaxDistance = 2500.0f;
    p->rayMarchingStepSize = 1.0f;
    p->lightSourceWorldPos[0] = -20.0f, p->lightSourceWorldPos[1] = 15.0f, p->lightSourceWorldPos[2] = -15.0f;
    p->beamRadius = 0.1f;
    p->lightRayStepSize = 0.3f;
    p->radiusFalloff = 0.0f;
}

int vr_set_grid(vr_ctx* c, const void* bytes, uint64_t size) {
    if (!c) return VR_E_ARG;
    if (!bytes) return fail(c, VR_E_ARG, "vr_set_grid: NULL buffer");
    const uint8_t* b = (const uint8_t*)bytes;
    if (size < 672 + 64 + 64) return fail(c, VR_E_BADGRID, "buffer too small for a NanoVDB grid (%llu bytes)", (unsigned long long)size);
    const uint64_t magic = rd<uint64_t>(b, 0);
    // "NanoVDB0" (grid), "NanoVDB1" / "NanoVDB2" are used by newer writers for the same GridData layout
    if (magic != 0x304244566f6e614eULL && magic != 0x314244566f6e614eULL && magic != 0x324244566f6e614eULL)
        return fail(c, VR_E_BADGRID, "bad NanoVDB magic 0x%016llx", (unsigned long long)magic);
    if (rd<uint32_t>(b, 636) != 1u) return fail(c, VR_E_BADGRID, "grid type %u is not FLOAT(1)", rd<uint32_t>(b, 636));
    if (size % 32 != 0) return fail(c, VR_E_BADGRID, "grid size %llu is not a multiple of 32 bytes", (unsigned long long)size);
    const uint64_t grid_size = rd<uint64_t>(b, 32);
    if (grid_size > size) return fail(c, VR_E_BADGRID, "gridSize field %llu exceeds the buffer (%llu)", (unsigned long long)grid_size, (unsigned long long)size);
    const uint64_t tree = 672;
    const uint64_t root = tree + rd<uint64_t>(b, tree + 24);
    if (root + 64 > size) return fail(c, VR_E_BADGRID, "root offset outside the buffer");
    const uint32_t tiles = rd<uint32_t>(b, root + 24);
    if (root + 64 + 32ull * tiles > size) return fail(c, VR_E_BADGRID, "root tile table outside the buffer");
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    if (c->grid && c->grid_cap < size) {
        cudaFree(c->grid);
        c->grid = nullptr;
        c->grid_cap = 0;
    }
    c->have_grid = false;
    if (!c->grid) {
        VR_CUDA(c, cudaMalloc(&c->grid, size));
        c->grid_cap = size;
    }
    VR_CUDA(c, cudaMemcpyAsync(c->grid, bytes, size, cudaMemcpyHostToDevice, c->stream));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    VrGridDev g{};
    g.base = c->grid;
    g.root = root;
    g.tile_count = tiles;
    for (int a = 0; a < 3; a++) {
        g.bmin[a] = rd<int32_t>(b, root + 0 + 4 * a);
        g.bmax[a] = rd<int32_t>(b, root + 12 + 4 * a);
        g.vec[a] = rd<float>(b, 296 + 72 + 4 * a);
    }
    for (int a = 0; a < 9; a++) {
        g.mat[a] = rd<float>(b, 296 + 0 + 4 * a);
        g.inv[a] = rd<float>(b, 296 + 36 + 4 * a);
    }
    c->gdev = g;
    c->grid_size = size;
    c->grid_gen++;
    c->have_grid = true;
    c->march_key.valid = false;
    if (size >= (1ull << 32)) c->err = "note: grid >= 4 GiB, beyond the reference's 32-bit GLSL addressing";
    return VR_OK;
}

int vr_resize(vr_ctx* c, uint32_t w, uint32_t h) {
    if (!c) return VR_E_ARG;
    if (w == 0 || h == 0 || w > 65536 || h > 65536) return fail(c, VR_E_ARG, "image size %ux%u invalid", w, h);
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    if (c->accum) cudaFree(c->accum), c->accum = nullptr;
    if (c->rgba8) cudaFree(c->rgba8), c->rgba8 = nullptr;
    if (c->frame_color) cudaFree(c->frame_color), c->frame_color = nullptr;
    const uint64_t npix = (uint64_t)w * h;
    VR_CUDA(c, cudaMalloc(&c->accum, npix * sizeof(float4)));
    VR_CUDA(c, cudaMalloc(&c->rgba8, npix * sizeof(uchar4)));
    VR_CUDA(c, cudaMalloc(&c->frame_color, npix * sizeof(float)));
    c->width = w, c->height = h;
    c->ubo.framebufferDim[0] = w, c->ubo.framebufferDim[1] = h;
    c->ubo.frameCount = 0;  // src/main.cpp:937
    c->march_key.valid = false;
    vrk_clear_image(c->stream, c->accum, c->rgba8, c->frame_color, npix);
    return check_stream(c);
}

int vr_set_params(vr_ctx* c, const vr_ubo* p) {
    if (!c) return VR_E_ARG;
    if (!p) return fail(c, VR_E_ARG, "vr_set_params: NULL params");
    // values the reference UI allows (src/main.cpp:662-691) but its shaders cannot survive
    if (!(p->rayMarchingStepSize > 0.0f)) return fail(c, VR_E_ARG, "rayMarchingStepSize must be > 0 (the shader loop never ends at 0, src/main.cpp:684)");
    if (p->maxLights > c->light_capacity)
        return fail(c, VR_E_CAPACITY, "maxLights %u exceeds the light buffer capacity %u (the reference writes out of bounds, src/main.cpp:676 vs :383-389)",
                    p->maxLights, c->light_capacity);
    if (c->width && (p->framebufferDim[0] != c->width || p->framebufferDim[1] != c->height))
        return fail(c, VR_E_ARG, "framebufferDim %ux%u does not match the image %ux%u (call vr_resize)", p->framebufferDim[0], p->framebufferDim[1], c->width, c->height);
    if (!(p->rayMaxDistance == p->rayMaxDistance) || !(p->fov == p->fov)) return fail(c, VR_E_ARG, "NaN parameter");
    c->ubo = *p;
    return VR_OK;
}

int vr_get_params(const vr_ctx* c, vr_ubo* p) {
    if (!c || !p) return VR_E_ARG;
    *p = c->ubo;
    return VR_OK;
}

int vr_set_max_lights_capacity(vr_ctx* c, uint32_t n) {
    if (!c) return VR_E_ARG;
    if (n == 0 || n > (1u << 28)) return fail(c, VR_E_ARG, "light capacity %u out of range", n);
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    c->lights.release();
    VR_CUDA(c, c->lights.reserve(n));
    VR_CUDA(c, cudaMemsetAsync(c->lights.p, 0, c->lights.cap * sizeof(vr_ray_light), c->stream));
    c->light_capacity = n;
    if (c->ubo.maxLights > n) c->ubo.maxLights = n;
    return vr_reset_lights(c);
}

int vr_reset_lights(vr_ctx* c) {
    if (!c) return VR_E_ARG;
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaMemsetAsync(c->counter, 0, sizeof(VrLightCounter), c->stream));
    return VR_OK;
}

int vr_write_lights(vr_ctx* c, const vr_ray_light* lights, uint32_t count) {
    if (!c) return VR_E_ARG;
    if (count > c->light_capacity) return fail(c, VR_E_CAPACITY, "%u lights exceed the capacity %u", count, c->light_capacity);
    if (count && !lights) return fail(c, VR_E_ARG, "vr_write_lights: NULL lights");
    VR_CUDA(c, cudaSetDevice(c->device));
    if (count) VR_CUDA(c, cudaMemcpyAsync(c->lights.p, lights, (size_t)count * sizeof(vr_ray_light), cudaMemcpyHostToDevice, c->stream));
    VrLightCounter h{count, 0u, (unsigned long long)count};
    VR_CUDA(c, cudaMemcpyAsync(c->counter, &h, sizeof(h), cudaMemcpyHostToDevice, c->stream));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    return VR_OK;
}

int vr_clear_image(vr_ctx* c) {
    if (!c) return VR_E_ARG;
    if (!c->accum) return fail(c, VR_E_STATE, "no image: call vr_resize first");
    VR_CUDA(c, cudaSetDevice(c->device));
    vrk_clear_image(c->stream, c->accum, c->rgba8, c->frame_color, (uint64_t)c->width * c->height);
    return check_stream(c);
}

int vr_dispatch_light_gen(vr_ctx* c, vr_algo algo, uint32_t gx, uint32_t gy, uint32_t gz, vr_lightgen_variant variant) {
    if (!c) return VR_E_ARG;
    if (!c->have_grid) return fail(c, VR_E_STATE, "no grid: call vr_set_grid first");
    if ((int)algo < 0 || (int)algo > 4) return fail(c, VR_E_ARG, "unknown algorithm %d", (int)algo);
    if (gx == 0 || gy == 0 || gz == 0) return fail(c, VR_E_ARG, "dispatch dimensions must be >= 1");
    const uint64_t nphot = (uint64_t)gx * 4 * gy * 4 * gz;
    if (nphot > (1ull << 28)) return fail(c, VR_E_ARG, "%llu photons exceed the supported dispatch size", (unsigned long long)nphot);
    if (c->ubo.maxLights > c->light_capacity) return fail(c, VR_E_CAPACITY, "maxLights %u exceeds capacity %u", c->ubo.maxLights, c->light_capacity);
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaEventRecord(c->ev[0], c->stream);
    c->lg_timed = true;
    if (algo != VR_ALGO_PATH) {  // path_light_gen.comp has an empty main (shaders/path_light_gen.comp:9-11)
        const uint32_t nx = gx * 4, ny = gy * 4, nz = gz;  // local_size 4x4x1 (shaders/light_gen.comp:7)
        VR_CUDA(c, c->photon_counts.reserve(nphot));
        VR_CUDA(c, c->photon_offsets.reserve(nphot));
        VR_CUDA(c, c->scan_tmp.reserve((nphot + 4095) / 4096 + 1));
        const int keep = (variant == VR_LG_KEEP_ORIGIN) ? 1 : 0;
        vrk_lightgen_count(c->stream, c->gdev, c->ubo, keep, nx, ny, nz, c->photon_counts.p);
        vrk_scan_u32(c->stream, c->photon_counts.p, c->photon_offsets.p, c->photon_total, nphot, c->scan_tmp.p);
        vrk_lightgen_emit(c->stream, c->gdev, c->ubo, keep, nx, ny, nz, c->photon_offsets.p, c->photon_total, c->lights.p, c->counter);
    }
    cudaEventRecord(c->ev[1], c->stream);
    return check_stream(c);
}

int vr_dispatch_compute_color(vr_ctx* c, vr_algo algo, const vr_tile* tile_or_null) {
    if (!c) return VR_E_ARG;
    if (!c->have_grid) return fail(c, VR_E_STATE, "no grid: call vr_set_grid first");
    if (!c->accum) return fail(c, VR_E_STATE, "no image: call vr_resize first");
    if ((int)algo < 0 || (int)algo > 4) return fail(c, VR_E_ARG, "unknown algorithm %d", (int)algo);
    if ((algo == VR_ALGO_RAY || algo == VR_ALGO_BEAM) && !(c->ubo.lightRayStepSize > 0.0f))
        return fail(c, VR_E_ARG, "lightRayStepSize must be > 0 for ray/beam (uint(len/0) is undefined, ray_compute_color.comp:15)");
    if (!(c->ubo.rayMarchingStepSize > 0.0f)) return fail(c, VR_E_ARG, "rayMarchingStepSize must be > 0");
    vr_tile tile;
    int rc = validate_tile(c, tile_or_null, &tile);
    if (rc != VR_OK) return rc;
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    const float tan_scale = vr_fov_scale(c->ubo.fov);
    const uint64_t npix = (uint64_t)tile.w * tile.h;
    memset(&c->last_counters, 0, sizeof(c->last_counters));

    if (algo == VR_ALGO_PATH) {
        cudaEventRecord(c->ev[2], st);
        cudaEventRecord(c->ev[3], st);
        cudaEventRecord(c->ev[4], st);
        cudaEventRecord(c->ev[5], st);
        VR_CUDA(c, cudaMemsetAsync(c->wc, 0, sizeof(VrWorkCounters), st));
        vrk_path(st, c->gdev, c->ubo, tan_scale, tile, c->frame_color, c->accum, c->rgba8, c->wc);
        cudaEventRecord(c->ev[6], st);
        VR_CUDA(c, cudaMemcpyAsync(c->wc_host, c->wc, sizeof(VrWorkCounters), cudaMemcpyDeviceToHost, st));
        cudaEventRecord(c->ev[7], st);
        c->timings_pending = true;
        VR_CUDA(c, cudaStreamSynchronize(st));
        c->last_counters.S_ref = c->wc_host->S_ref;
        c->last_counters.S_eff = c->wc_host->S_eff;
        c->last_counters.E = c->wc_host->S_eff;
        return check_stream(c);
    }

    // ---- light list for the evaluation kernel
    cudaEventRecord(c->ev[2], st);
    const bool expand = (algo == VR_ALGO_RAY || algo == VR_ALGO_BEAM);
    // lightCount lives on the device; launches that must not wait for it are sized by the buffer capacity
    const uint32_t nl_launch = c->light_capacity;
    if (expand) {
        VR_CUDA(c, c->light_steps.reserve(nl_launch));
        VR_CUDA(c, c->light_offsets.reserve(nl_launch));
        VR_CUDA(c, c->scan_tmp.reserve((nl_launch + 4095) / 4096 + 1));
        vrk_prepare_count_steps(st, c->lights.p, c->counter, c->ubo.lightRayStepSize, c->light_steps.p, nl_launch);
        vrk_scan_u32(st, c->light_steps.p, c->light_offsets.p, c->totals + 1, nl_launch, c->scan_tmp.p);
    }
    cudaEventRecord(c->ev[3], st);

    // ---- camera march: count -> scan (or reuse)
    MarchKey key;
    key.grid_gen = c->grid_gen, key.tile = tile;
    key.dim[0] = c->width, key.dim[1] = c->height;
    memcpy(key.cam, c->ubo.cameraPos, sizeof(key.cam));
    key.fov = c->ubo.fov, key.maxd = c->ubo.rayMaxDistance, key.step = c->ubo.rayMarchingStepSize, key.absorb = c->ubo.absorptionCoefficient;
    key.valid = true;
    const bool reuse = c->march_cache && (key == c->march_key);
    if (!reuse) {
        c->march_key.valid = false;
        VR_CUDA(c, c->px_counts.reserve(npix));
        VR_CUDA(c, c->px_offsets.reserve(npix));
        VR_CUDA(c, c->scan_tmp.reserve((npix + 4095) / 4096 + 1));
        VR_CUDA(c, cudaMemsetAsync(c->wc, 0, sizeof(VrWorkCounters), st));
        vrk_march(st, c->gdev, c->ubo, tan_scale, tile, 0, c->px_counts.p, nullptr, nullptr, c->wc);
        vrk_scan_u32(st, c->px_counts.p, c->px_offsets.p, c->totals + 0, npix, c->scan_tmp.p);
        VR_CUDA(c, cudaMemcpyAsync(c->wc_host, c->wc, sizeof(VrWorkCounters), cudaMemcpyDeviceToHost, st));
    }
    cudaEventRecord(c->ev[4], st);
    // one host round trip per frame: the two totals size the launches and buffers below
    VR_CUDA(c, cudaMemcpyAsync(c->totals_host, c->totals, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    VR_CUDA(c, cudaMemcpyAsync(c->totals_host + 2, &c->counter->lightCount, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    VR_CUDA(c, cudaStreamSynchronize(st));
    uint32_t light_count_host = (uint32_t)(c->totals_host[2] & 0xFFFFFFFFull);
    rc = check_stream(c);
    if (rc != VR_OK) return rc;
    const uint64_t nsamples = reuse ? c->cached_samples : c->totals_host[0];
    if (light_count_host > c->light_capacity) light_count_host = c->light_capacity;
    const uint64_t neval = expand ? c->totals_host[1] : (uint64_t)light_count_host;
    if (neval > (1ull << 32)) return fail(c, VR_E_ARG, "expanded light list of %llu entries is too large (lightRayStepSize too small?)", (unsigned long long)neval);
    VR_CUDA(c, c->eval_lights.reserve(neval + 1));
    if (expand)
        vrk_prepare_expand(st, c->lights.p, c->counter, c->ubo.lightRayStepSize, c->light_offsets.p, c->eval_lights.p, light_count_host, neval);
    else
        vrk_prepare_simple(st, c->lights.p, c->counter, algo == VR_ALGO_SPHERE ? 1 : 0, c->eval_lights.p, light_count_host);
    if (!reuse) {
        VR_CUDA(c, c->samples.reserve(nsamples + 1));
        vrk_march(st, c->gdev, c->ubo, tan_scale, tile, 1, c->px_counts.p, c->px_offsets.p, c->samples.p, nullptr);
        c->cached_samples = nsamples;
        c->march_key = key;
        c->last_counters.S_ref = c->wc_host->S_ref;
        c->last_counters.S_eff = c->wc_host->S_eff;
    } else {
        c->last_counters.S_ref = c->wc_host->S_ref;
        c->last_counters.S_eff = c->wc_host->S_eff;
    }
    VR_CUDA(c, c->contrib.reserve(nsamples + 1));
    cudaEventRecord(c->ev[5], st);
    // ---- evaluation (hot) + resolve
    const int sphere = (algo == VR_ALGO_SPHERE || algo == VR_ALGO_BEAM) ? 1 : 0;
    if (!sphere && c->eval_variant >= 1) {
        VR_CUDA(c, c->eval_lights2.reserve(2 * neval + 10));
        vrk_eval_point_packed(st, c->eval_variant, c->samples.p, nsamples, c->eval_lights.p, c->eval_lights2.p, neval, c->contrib.p,
                              c->eval_redo);
    } else if (sphere && c->eval_variant >= 1) {
        VR_CUDA(c, c->eval_lights2.reserve(2 * neval + 10));
        vrk_eval_sphere_packed(st, c->samples.p, nsamples, c->eval_lights.p, c->eval_lights2.p, neval, c->ubo.beamRadius, c->contrib.p);
    } else {
        vrk_eval(st, sphere, c->samples.p, nsamples, c->eval_lights.p, neval, c->ubo.beamRadius, c->contrib.p);
    }
    cudaEventRecord(c->ev[6], st);
    vrk_resolve(st, c->ubo, tile, c->px_counts.p, c->px_offsets.p, c->contrib.p, c->counter, c->frame_color, c->accum, c->rgba8);
    cudaEventRecord(c->ev[7], st);
    c->timings_pending = true;
    c->last_counters.eval_lights = neval;
    c->last_counters.light_count = light_count_host;
    c->last_counters.E = nsamples * neval;
    return check_stream(c);
}

int vr_set_light_gen_groups(vr_ctx* c, uint32_t gx, uint32_t gy, uint32_t gz, vr_lightgen_variant variant) {
    if (!c) return VR_E_ARG;
    if (gx == 0 || gy == 0 || gz == 0) return fail(c, VR_E_ARG, "dispatch dimensions must be >= 1");
    c->lg_groups[0] = gx, c->lg_groups[1] = gy, c->lg_groups[2] = gz;
    c->lg_variant = variant;
    return VR_OK;
}

int vr_render_frame(vr_ctx* c, vr_algo algo, const vr_tile* tile_or_null) {
    if (!c) return VR_E_ARG;
    if (!c->have_grid) return fail(c, VR_E_STATE, "no grid: call vr_set_grid first");
    if (!c->accum) return fail(c, VR_E_STATE, "no image: call vr_resize first");
    c->ubo.frameCount += 1;  // src/main.cpp:712
    int rc = vr_reset_lights(c);  // :722-728
    if (rc != VR_OK) return rc;
    if (c->ubo.frameCount == 1) {  // :731
        rc = vr_clear_image(c);
        if (rc != VR_OK) return rc;
    }
    rc = vr_dispatch_light_gen(c, algo, c->lg_groups[0], c->lg_groups[1], c->lg_groups[2], c->lg_variant);  // :811-814
    if (rc != VR_OK) return rc;
    return vr_dispatch_compute_color(c, algo, tile_or_null);  // :832-837
}

int vr_reset_accumulation(vr_ctx* c) {
    if (!c) return VR_E_ARG;
    c->ubo.frameCount = 0;
    return VR_OK;
}

int vr_sync(vr_ctx* c) {
    if (!c) return VR_E_ARG;
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    return check_stream(c);
}

int vr_read_light_counter(vr_ctx* c, uint32_t counter[2], uint64_t* total_emitted) {
    if (!c) return VR_E_ARG;
    VR_CUDA(c, cudaSetDevice(c->device));
    VrLightCounter h;
    VR_CUDA(c, cudaMemcpyAsync(&h, c->counter, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    if (counter) counter[0] = h.lightCount, counter[1] = h.contributionCounter;
    if (total_emitted) *total_emitted = h.totalEmitted;
    return VR_OK;
}

int vr_read_lights(vr_ctx* c, vr_ray_light* out, uint32_t capacity, uint32_t* count) {
    if (!c) return VR_E_ARG;
    uint32_t cnt[2];
    int rc = vr_read_light_counter(c, cnt, nullptr);
    if (rc != VR_OK) return rc;
    if (count) *count = cnt[0];
    uint32_t n = cnt[0] < c->light_capacity ? cnt[0] : c->light_capacity;
    if (n > capacity) n = capacity;
    if (n && out) {
        VR_CUDA(c, cudaMemcpyAsync(out, c->lights.p, (size_t)n * sizeof(vr_ray_light), cudaMemcpyDeviceToHost, c->stream));
        VR_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return VR_OK;
}

#define VR_READ_IMAGE(c, dst, src, bytes)                                                           \
    do {                                                                                            \
        if (!(c)) return VR_E_ARG;                                                                  \
        if (!(c)->accum) return fail((c), VR_E_STATE, "no image: call vr_resize first");            \
        if (!(dst)) return fail((c), VR_E_ARG, "NULL output buffer");                               \
        VR_CUDA((c), cudaSetDevice((c)->device));                                                   \
        VR_CUDA((c), cudaMemcpyAsync((dst), (src), (bytes), cudaMemcpyDeviceToHost, (c)->stream));  \
        VR_CUDA((c), cudaStreamSynchronize((c)->stream));                                           \
        return VR_OK;                                                                               \
    } while (0)

int vr_read_image_rgba8(vr_ctx* c, uint8_t* out) { VR_READ_IMAGE(c, out, c->rgba8, (size_t)c->width * c->height * 4); }
int vr_read_accum_f32(vr_ctx* c, float* out) { VR_READ_IMAGE(c, out, c->accum, (size_t)c->width * c->height * 16); }
int vr_read_frame_color_f32(vr_ctx* c, float* out) { VR_READ_IMAGE(c, out, c->frame_color, (size_t)c->width * c->height * 4); }

int vr_write_accum_f32(vr_ctx* c, const float* in, uint32_t frame_count) {
    if (!c) return VR_E_ARG;
    if (!c->accum) return fail(c, VR_E_STATE, "no image: call vr_resize first");
    if (!in) return fail(c, VR_E_ARG, "NULL input buffer");
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaMemcpyAsync(c->accum, in, (size_t)c->width * c->height * 16, cudaMemcpyHostToDevice, c->stream));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    c->ubo.frameCount = frame_count;
    return VR_OK;
}
int vr_write_image_rgba8(vr_ctx* c, const uint8_t* in) {
    if (!c) return VR_E_ARG;
    if (!c->accum) return fail(c, VR_E_STATE, "no image: call vr_resize first");
    if (!in) return fail(c, VR_E_ARG, "NULL input buffer");
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaMemcpyAsync(c->rgba8, in, (size_t)c->width * c->height * 4, cudaMemcpyHostToDevice, c->stream));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    return VR_OK;
}

int vr_get_counters(vr_ctx* c, vr_counters* out) {
    if (!c || !out) return VR_E_ARG;
    *out = c->last_counters;
    return VR_OK;
}

int vr_get_timings(vr_ctx* c, vr_timings* out) {
    if (!c || !out) return VR_E_ARG;
    VR_CUDA(c, cudaSetDevice(c->device));
    int rc = finalize_timings(c);
    if (rc != VR_OK) return rc;
    *out = c->timings;
    return VR_OK;
}

int vr_get_device_buffers(vr_ctx* c, void** accum_f32, void** image_rgba8, void** frame_color_f32, void** stream) {
    if (!c) return VR_E_ARG;
    if (accum_f32) *accum_f32 = c->accum;
    if (image_rgba8) *image_rgba8 = c->rgba8;
    if (frame_color_f32) *frame_color_f32 = c->frame_color;
    if (stream) *stream = (void*)c->stream;
    return VR_OK;
}

int vr_row_sample_counts(vr_ctx* c, uint64_t* rows_out) {
    if (!c) return VR_E_ARG;
    if (!c->have_grid) return fail(c, VR_E_STATE, "no grid: call vr_set_grid first");
    if (!c->accum) return fail(c, VR_E_STATE, "no image: call vr_resize first");
    if (!rows_out) return fail(c, VR_E_ARG, "NULL output buffer");
    VR_CUDA(c, cudaSetDevice(c->device));
    const vr_tile tile = full_tile(c);
    const uint64_t npix = (uint64_t)tile.w * tile.h;
    c->march_key.valid = false;
    VR_CUDA(c, c->px_counts.reserve(npix));
    DevBuf<unsigned long long> rows;
    VR_CUDA(c, rows.reserve(c->height));
    vrk_march(c->stream, c->gdev, c->ubo, vr_fov_scale(c->ubo.fov), tile, 0, c->px_counts.p, nullptr, nullptr, nullptr);
    vrk_row_sums(c->stream, c->px_counts.p, c->width, c->height, rows.p);
    cudaError_t e = cudaMemcpyAsync(rows_out, rows.p, (size_t)c->height * 8, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    rows.release();
    if (e != cudaSuccess) return fail(c, VR_E_CUDA, "row sample counts failed: %s", cudaGetErrorString(e));
    return check_stream(c);
}

int vr_set_stream(vr_ctx* c, void* cuda_stream) {
    if (!c) return VR_E_ARG;
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
    return VR_OK;
}
int vr_get_launch_count(vr_ctx* c, uint64_t* launches) {
    if (!c || !launches) return VR_E_ARG;
    *launches = vrk_launch_count() - c->launch_base;
    return VR_OK;
}

int vr_timer_start(vr_ctx* c) {
    if (!c) return VR_E_ARG;
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    VR_CUDA(c, cudaEventRecord(c->tev[0], c->stream));
    return VR_OK;
}
int vr_timer_stop(vr_ctx* c, float* ms) {
    if (!c || !ms) return VR_E_ARG;
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaEventRecord(c->tev[1], c->stream));
    VR_CUDA(c, cudaEventSynchronize(c->tev[1]));
    VR_CUDA(c, cudaEventElapsedTime(ms, c->tev[0], c->tev[1]));
    return VR_OK;
}
int vr_ubench(vr_ctx* c, int kind, double* result) {
    if (!c || !result || kind < 0 || kind > 4) return VR_E_ARG;
    VR_CUDA(c, cudaSetDevice(c->device));
    float* sink = nullptr;
    VR_CUDA(c, cudaMalloc(&sink, 64));
    vrk_ubench(c->stream, kind, sink, 64);  // warm-up
    double best = 0.0;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(c->tev[0], c->stream);
        const double ops = vrk_ubench(c->stream, kind, sink, 4096);
        cudaEventRecord(c->tev[1], c->stream);
        cudaError_t e = cudaEventSynchronize(c->tev[1]);
        if (e != cudaSuccess) {
            cudaFree(sink);
            return fail(c, VR_E_CUDA, "ubench failed: %s", cudaGetErrorString(e));
        }
        float ms = 0;
        cudaEventElapsedTime(&ms, c->tev[0], c->tev[1]);
        const double rate = ops / (ms * 1e-3) * (kind == 0 ? 2.0 : (kind == 3 ? 4.0 : 1.0)) / 1e12;
        if (rate > best) best = rate;
    }
    cudaFree(sink);
    *result = best;
    return check_stream(c);
}

int vr_lookup_values(vr_ctx* c, const int32_t* ijk, uint64_t n, float* values_out) {
    if (!c) return VR_E_ARG;
    if (!c->have_grid) return fail(c, VR_E_STATE, "no grid: call vr_set_grid first");
    if (n == 0) return VR_OK;
    if (!ijk || !values_out) return fail(c, VR_E_ARG, "vr_lookup_values: NULL buffer");
    VR_CUDA(c, cudaSetDevice(c->device));
    int32_t* d_ijk = nullptr;
    float* d_out = nullptr;
    VR_CUDA(c, cudaMalloc(&d_ijk, n * 12));
    cudaError_t e = cudaMalloc(&d_out, n * 4);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_ijk, ijk, n * 12, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) {
        vrk_lookup(c->stream, c->gdev, d_ijk, n, d_out);
        e = cudaMemcpyAsync(values_out, d_out, n * 4, cudaMemcpyDeviceToHost, c->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d_ijk);
    if (d_out) cudaFree(d_out);
    if (e != cudaSuccess) return fail(c, VR_E_CUDA, "vr_lookup_values failed: %s", cudaGetErrorString(e));
    return check_stream(c);
}

int vr_set_eval_variant(vr_ctx* c, int variant) {
    if (!c || variant < 0 || variant > 3) return VR_E_ARG;
    c->eval_variant = variant;
    return VR_OK;
}

int vr_read_samples(vr_ctx* c, float* out, uint64_t capacity, uint64_t* count) {
    if (!c) return VR_E_ARG;
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    const uint64_t n = c->cached_samples;
    if (count) *count = n;
    const uint64_t m = n < capacity ? n : capacity;
    if (m && !out) return fail(c, VR_E_ARG, "vr_read_samples: NULL output buffer");
    if (m) {
        VR_CUDA(c, cudaMemcpyAsync(out, c->samples.p, m * sizeof(VrSample), cudaMemcpyDeviceToHost, c->stream));
        VR_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return VR_OK;
}

int vr_get_eval_redo_count(vr_ctx* c, uint64_t* ctas) {
    if (!c || !ctas) return VR_E_ARG;
    VR_CUDA(c, cudaSetDevice(c->device));
    unsigned long long h = 0;
    VR_CUDA(c, cudaMemcpyAsync(&h, c->eval_redo, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    *ctas = h;
    return VR_OK;
}

int vr_set_march_cache(vr_ctx* c, int enabled) {
    if (!c) return VR_E_ARG;
    c->march_cache = enabled != 0;
    c->march_key.valid = false;
    return VR_OK;
}

int vr_cpu_test_render(vr_ctx* c, uint32_t width, uint32_t height, float* density_out) {
    if (!c) return VR_E_ARG;
    if (!c->have_grid) return fail(c, VR_E_STATE, "no grid: call vr_set_grid first");
    if (width == 0 || height == 0 || !density_out) return fail(c, VR_E_ARG, "vr_cpu_test_render: bad arguments");
    VR_CUDA(c, cudaSetDevice(c->device));
    float* d = nullptr;
    VR_CUDA(c, cudaMalloc(&d, (size_t)width * height * 4));
    // CPU_test/main.cpp:99-100: aspect in float, scale = tan(fov*0.5*M_PI/180.0) in double narrowed to float
    const float fov = 45.0;
    const float aspect = static_cast<float>(width) / height;
    const float scale = (float)tan(fov * 0.5 * M_PI / 180.0);
    cudaEventRecord(c->ev[5], c->stream);
    vrk_cpu_test(c->stream, c->gdev, width, height, aspect, scale, d);
    cudaEventRecord(c->ev[6], c->stream);
    cudaError_t e = cudaMemcpyAsync(density_out, d, (size_t)width * height * 4, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev[5], c->ev[6]);
    memset(&c->timings, 0, sizeof(c->timings));
    c->timings.eval_ms = ms, c->timings.total_ms = ms;
    c->timings_pending = false;
    cudaFree(d);
    if (e != cudaSuccess) return fail(c, VR_E_CUDA, "cpu_test render failed: %s", cudaGetErrorString(e));
    return check_stream(c);
}

}  // extern "C"
