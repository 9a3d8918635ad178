// vr_api.cu -- the C ABI of include/vr.h: a headless context that plays the role of the descriptor set + command
// buffer of VolumeApp::drawFrame (src/main.cpp:712-837) on one CUDA stream. No exceptions cross the ABI.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <new>
#include <string>

#include "../../include/vr.h"
#include "vr_kernels.cuh"

// NVTX ranges per pass of a frame (tracing row), compiled in with -DVR_NVTX (VR_NVTX=1 in the
// environment of __graft_entry__.build(); header-only NVTX 3, no link dependency). A VrPassRange names the pass that
// is being ENQUEUED; profilers project the range onto the stream's kernels. Off by default: the product build carries
// no profiler hooks.
#ifdef VR_NVTX
#include <nvtx3/nvToolsExt.h>
struct VrPassRange {
    bool open = false;
    explicit VrPassRange(const char* name) { next(name); }
    void next(const char* name) {
        if (open) nvtxRangePop();
        nvtxRangePushA(name);
        open = true;
    }
    ~VrPassRange() {
        if (open) nvtxRangePop();
    }
};
#else
struct VrPassRange {
    explicit VrPassRange(const char*) {}
    void next(const char*) {}
};
#endif

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
    uint32_t trilinear = 0;
    uint32_t gaps = 0;  // layout of the sample buffer: 1 = single-pass march, one slot per pixel sized by its step bound
    bool valid = false;
    bool operator==(const MarchKey& o) const {
        return valid && o.valid && grid_gen == o.grid_gen && memcmp(&tile, &o.tile, sizeof(tile)) == 0 &&
               memcmp(dim, o.dim, sizeof(dim)) == 0 && memcmp(cam, o.cam, sizeof(cam)) == 0 && fov == o.fov &&
               maxd == o.maxd && step == o.step && absorb == o.absorb && trilinear == o.trilinear && gaps == o.gaps;
    }
};
}  // namespace

// what the per-pixel step bounds of the single-pass march depend on: the rays and the grid's box and map, not its voxels
struct BoundKey {
    vr_tile tile = {0, 0, 0, 0};
    uint32_t dim[2] = {0, 0};
    float cam[3] = {0, 0, 0}, fov = 0, maxd = 0, step = 0;
    int32_t bmin[3] = {0, 0, 0}, bmax[3] = {0, 0, 0};
    float mat[9] = {0}, inv[9] = {0}, vec[3] = {0};
    bool valid = false;
    bool operator==(const BoundKey& o) const {
        return valid && o.valid && memcmp(&tile, &o.tile, sizeof(tile)) == 0 && memcmp(dim, o.dim, sizeof(dim)) == 0 &&
               memcmp(cam, o.cam, sizeof(cam)) == 0 && fov == o.fov && maxd == o.maxd && step == o.step &&
               memcmp(bmin, o.bmin, sizeof(bmin)) == 0 && memcmp(bmax, o.bmax, sizeof(bmax)) == 0 &&
               memcmp(mat, o.mat, sizeof(mat)) == 0 && memcmp(inv, o.inv, sizeof(inv)) == 0 && memcmp(vec, o.vec, sizeof(vec)) == 0;
    }
};

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
    // second light buffer + counter: while a frame is being evaluated, vr_render_frame already generates the NEXT
    // frame's lights into these (same UBO, frameCount + 1) on the auxiliary stream; the next call adopts them by
    // swapping the pairs if its inputs are what was assumed (LgSpec), else it regenerates.
    DevBuf<vr_ray_light> lights_alt;
    VrLightCounter* counter_alt = nullptr;
    struct LgSpec {
        bool valid = false;
        vr_ubo ubo{};
        uint64_t grid_gen = 0;
        uint32_t groups[3] = {0, 0, 0};
        int variant = 0;
    } spec;
    vr_ubo prev_frame_ubo{};       // UBO of the previous vr_render_frame (steady-state detection)
    uint64_t prev_frame_grid_gen = 0;
    bool prev_frame_valid = false;
    bool lg_prefetch = true, prefetch_next = false;
    DevBuf<uint32_t> photon_counts;
    DevBuf<unsigned long long> photon_offsets;
    DevBuf<unsigned long long> lg_scan_tmp;  // light generation runs on the auxiliary stream: own scan scratch
    DevBuf<float4> lg_slab;                  // per-photon private light lists (vrk_lightgen)
    unsigned long long* photon_total = nullptr;
    bool lights_external = false;            // list came from vr_write_lights: no bound on the segment lengths
    uint32_t external_count = 0;
    // light generation of vr_render_frame overlaps the camera march on this stream
    cudaStream_t aux_stream = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_lg_done = nullptr;
    // ev_frame_end[p]: end of the last vr_render_frame that used light pair parity p (its last reader of lights/counter);
    // ev_grid_ready: the bytes of the installed grid have arrived (vr_set_grid_device*: a copy that may still be in flight)
    cudaEvent_t ev_frame_end[2] = {nullptr, nullptr}, ev_grid_ready = nullptr;
    bool frame_end_recorded[2] = {false, false}, grid_ready_pending = false;
    uint32_t pair_parity = 0;  // toggles with every swap of (lights, counter) <-> (lights_alt, counter_alt)
    cudaStream_t upload_stream = nullptr;  // vr_set_grid: host->device copy of a new grid next to the frames in flight
    uint8_t* grid_alt = nullptr;           // the second grid buffer (vr_set_grid alternates)
    uint64_t grid_alt_cap = 0;
    bool grid_alt_busy = false;
    // the buffer that is not the installed grid is free for the next grid once the work that was enqueued before the
    // last install has passed: ev_grid_alt_free on the render stream, ev_grid_alt_free_aux on the light-generation stream
    cudaEvent_t ev_grid_alt_free = nullptr, ev_grid_alt_free_aux = nullptr;
    cudaStream_t copy_stream = nullptr;  // asynchronous read-back (vr_read_accum_f32_async): copies overlap the next frame
    cudaEvent_t ev_copy_ready = nullptr, ev_copy_done = nullptr;
    DevBuf<float> accum_snapshot;
    bool copy_pending = false;
    cudaStream_t seg_stream2 = nullptr;
    cudaStream_t seg_stream = nullptr;  // low priority: the small evaluation kernels that fill the big kernel's tail
    cudaEvent_t ev_seg_fork = nullptr, ev_seg_join = nullptr, ev_seg_join2 = nullptr;
    bool lg_in_flight = false;
    uint32_t lg_groups[3] = {1, 1, 1};
    vr_lightgen_variant lg_variant = VR_LG_RESET_ORIGIN;
    // binding 0 (+ float accumulator and last frame colour)
    uint32_t width = 0, height = 0;
    float4* accum = nullptr;
    uchar4* rgba8 = nullptr;
    float* frame_color = nullptr;        // the buffer in use: own_frame_color or a caller-provided (symmetric) one
    float* own_frame_color = nullptr;
    VrFrameColorOut fc_out{};            // frame_color + peer mappings for the kernels
    // gather work buffers
    DevBuf<uint32_t> px_counts;
    DevBuf<unsigned long long> px_offsets;
    DevBuf<unsigned long long> scan_tmp;
    DevBuf<VrSample> samples;
    DevBuf<float> contrib;
    // ray-segment evaluation (eval_variant 8): segment table + index list of the exact kernel, rebuilt with the march
    DevBuf<uint32_t> seg_first, seg_pix, seg_flag, seg16_first, seg16_pix, seg16_flag, seg_list, seg_list2, seg_base, seg_sched;
    DevBuf<float> seg_part16, seg_part8;  // scratch of the slice units of k_eval_seg
    uint64_t seg_reserved_for = 0, seg_reserved_npix = 0;  // sample count / pixels the tables were last sized for
    DevBuf<uint8_t> seg_flags;
    unsigned long long* seg_counts = nullptr;  // device: [0] segments, [1] leftovers, [2] exact-list length
    bool seg_valid = false;                    // tables match the cached march
    int seg_pairs = -1;                        // ... and were built for point-type (1) / sphere-type (0) lists
    VrPhysics phys = {0, 0, 0, 0.0f};          // opt-in physics (vr_set_options); all zero = the reference's arithmetic
    float eval_near = 0.25f;                   // distance below which a (sample, light) term is left to the exact kernel
    DevBuf<VrEvalLight> eval_lights;
    DevBuf<float4> eval_lights2;
    DevBuf<float> path_tab;  // path mode: shadow-term table (k_path_table)
    int eval_variant = 8;  // 0: scalar kernel, 1: packed FP32x2 kernel, 2/3: packed speculative kernel, 8: ray-segment kernel (point/ray lists)
    DevBuf<uint32_t> light_steps;
    DevBuf<unsigned long long> light_offsets;
    unsigned long long* totals = nullptr;       // [0] samples, [1] expanded lights (device)
    unsigned long long* totals_host = nullptr;  // pinned: blocking read-backs on the slow paths
    VrWorkCounters* wc = nullptr;
    uint32_t* expand_overflow = nullptr;        // device flag: the expanded light list did not fit its buffer
    // per-frame scalars copied to pinned memory without blocking; vr_get_counters waits for ev_scalars only
    struct Scalars {
        unsigned long long nsamples, neval, S_ref, S_eff;
        uint32_t light_count, overflow;
    };
    Scalars* scalars_host = nullptr;
    cudaEvent_t ev_scalars = nullptr;
    bool counters_pending = false, counters_have_march = false;
    uint64_t nsamples_used = 0;  // what totals[0] holds for this gather (sample buffer range)
    uint64_t nreal_used = 0;     // samples of this gather
    unsigned long long* eval_redo = nullptr;  // CTAs of the speculative eval kernel that repeated with the guard (cumulative)
    vr_counters last_counters{};
    vr_algo last_algo = VR_ALGO_POINT;
    bool march_cache = false;
    MarchKey march_key;
    uint64_t cached_samples = 0;  // length of the sample buffer's address range (= sample count when the buffer is compact)
    uint64_t cached_real = 0;     // number of samples in it, or an upper bound of it (real_exact): sizes tables and launches
    bool real_exact = false;      // cached_real is the count itself
    bool cached_S_valid = false;  // cached_S holds the counters of the cached march
    bool march_single = true;     // vr_set_march_single_pass
    uint32_t frame_step = 1;      // vr_set_frame_step: frameCount advances by this much per vr_render_frame
    uint32_t frame_first = 0;     // ... after the next frame, which is number frame_first (0: no override)
    int accumulate_mode = 0;      // vr_set_accumulate_mode: 0 = running average (the reference), 1 = float sum
    bool eval_side_stream = true; // vr_set_eval_side_stream
    // what the current light list was generated from (vr_dispatch_light_gen / vr_render_frame): bounds the list length and
    // the ray/beam segment lengths without reading the device, whatever vr_set_params did since
    uint32_t lights_bound = 0;        // upper bound of counter->lightCount
    float gen_light_pos[3] = {0, 0, 0};
    int gen_pos_state = 0;            // 0: nothing generated since the counter reset, 1: every generation used gen_light_pos,
                                      // 2: unknown end points (lights written by the caller, or several source positions)
    MarchKey gaps_refused;        // rays whose bound total does not fit 32-bit sample indices: two-pass march
    BoundKey bound_key;           // px_offsets hold the slots of these rays (single-pass layout), bound_extent their total
    uint64_t bound_extent = 0;
    uint64_t cached_S[2] = {0, 0};  // S_ref, S_eff of the cached march
    // timing: a ring of event sets so that per-pass times can be collected without a host sync per frame
    struct TimingSet {
        cudaEvent_t ev[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
        // [0],[1] light generation (its own stream) . [2] gather start . [3] march done . [4] list ready . [5] eval done . [6] resolve done
        bool pending = false, lg_timed = false, open = false;
    };
    static constexpr int kTimingSets = 4;
    TimingSet tset[kTimingSets];
    int tcur = 0;
    cudaEvent_t tev[2] = {nullptr, nullptr};
    vr_timings timings{}, timings_sum{};
    uint64_t timings_frames = 0;
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

// ---- timing ring. A frame (or a lone dispatch) owns one TimingSet; its events are read when the set is reused or when
// the caller asks, so collecting per-pass times costs no host synchronisation per frame.
int finalize_set(vr_ctx* c, vr_ctx::TimingSet& t) {
    if (!t.pending) return VR_OK;
    VR_CUDA(c, cudaEventSynchronize(t.ev[6]));
    if (t.lg_timed) VR_CUDA(c, cudaEventSynchronize(t.ev[1]));
    vr_timings m{};
    float ms = 0;
    if (t.lg_timed) {
        cudaEventElapsedTime(&ms, t.ev[0], t.ev[1]);
        m.light_gen_ms = ms;
    }
    cudaEventElapsedTime(&ms, t.ev[2], t.ev[3]);
    m.march_ms = ms;
    cudaEventElapsedTime(&ms, t.ev[3], t.ev[4]);
    m.prepare_ms = ms;
    cudaEventElapsedTime(&ms, t.ev[4], t.ev[5]);
    m.eval_ms = ms;
    cudaEventElapsedTime(&ms, t.ev[5], t.ev[6]);
    m.resolve_ms = ms;
    cudaEventElapsedTime(&ms, t.ev[2], t.ev[6]);
    m.total_ms = ms;
    c->timings = m;
    c->timings_sum.light_gen_ms += m.light_gen_ms, c->timings_sum.march_ms += m.march_ms;
    c->timings_sum.prepare_ms += m.prepare_ms, c->timings_sum.eval_ms += m.eval_ms;
    c->timings_sum.resolve_ms += m.resolve_ms, c->timings_sum.total_ms += m.total_ms;
    c->timings_frames++;
    t.pending = t.lg_timed = t.open = false;
    return VR_OK;
}
int finalize_timings(vr_ctx* c) {
    for (int k = 1; k <= vr_ctx::kTimingSets; k++) {  // oldest first
        int rc = finalize_set(c, c->tset[(c->tcur + k) % vr_ctx::kTimingSets]);
        if (rc != VR_OK) return rc;
    }
    return VR_OK;
}
// the set the next dispatch records into: the open one (light generation already recorded), else a fresh one
int open_timing_set(vr_ctx* c, vr_ctx::TimingSet** out) {
    vr_ctx::TimingSet* t = &c->tset[c->tcur];
    if (!t->open) {
        c->tcur = (c->tcur + 1) % vr_ctx::kTimingSets;
        t = &c->tset[c->tcur];
        int rc = finalize_set(c, *t);
        if (rc != VR_OK) return rc;
        t->open = true;
    }
    *out = t;
    return VR_OK;
}

// Upper bound of the expanded ray/beam list, so that its buffer can be sized without reading the device: every light
// end point is the light source or a scatter position inside the root bbox (light_gen.comp:74-84), hence no segment is
// longer than the diagonal of the world-space box around both. 0 = no usable bound.
uint64_t expand_bound(const vr_ctx* c) {
    if (c->gen_pos_state != 1) return 0;
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    const VrGridDev& g = c->gdev;
    for (int corner = 0; corner < 8; corner++) {
        const double i = (corner & 1) ? (double)g.bmax[0] + 1.0 : (double)g.bmin[0];
        const double j = (corner & 2) ? (double)g.bmax[1] + 1.0 : (double)g.bmin[1];
        const double k = (corner & 4) ? (double)g.bmax[2] + 1.0 : (double)g.bmin[2];
        for (int a = 0; a < 3; a++) {
            const double w = i * g.mat[3 * a] + j * g.mat[3 * a + 1] + k * g.mat[3 * a + 2] + g.vec[a];
            lo[a] = w < lo[a] ? w : lo[a], hi[a] = w > hi[a] ? w : hi[a];
        }
    }
    for (int a = 0; a < 3; a++) {  // the source position the list was generated with, not the current UBO's
        const double w = c->gen_light_pos[a];
        lo[a] = w < lo[a] ? w : lo[a], hi[a] = w > hi[a] ? w : hi[a];
    }
    const double diag = sqrt((hi[0] - lo[0]) * (hi[0] - lo[0]) + (hi[1] - lo[1]) * (hi[1] - lo[1]) + (hi[2] - lo[2]) * (hi[2] - lo[2]));
    const double steps = floor(diag * 1.001 / (double)c->ubo.lightRayStepSize) + 2.0;
    const double bound = steps * (double)(c->lights_bound < c->light_capacity ? c->lights_bound : c->light_capacity);
    if (!(bound == bound) || bound > (double)(1ull << 26)) return 0;
    return (uint64_t)bound;
}

// bookkeeping for a light generation with parameters `u` into the CURRENT (lights, counter) pair: the shader appends at
// most up to index maxLights - 1 and stores lightCount = maxLights beyond (light_gen.comp:79-87), so the counter never
// exceeds the largest maxLights used since its reset
void note_generation(vr_ctx* c, const vr_ubo& u, vr_algo algo) {
    if (algo == VR_ALGO_PATH) return;
    if (u.maxLights > c->lights_bound) c->lights_bound = u.maxLights;
    if (c->gen_pos_state == 0) {
        memcpy(c->gen_light_pos, u.lightSourceWorldPos, sizeof(c->gen_light_pos));
        c->gen_pos_state = 1;
    } else if (c->gen_pos_state == 1 && memcmp(c->gen_light_pos, u.lightSourceWorldPos, sizeof(c->gen_light_pos)) != 0) {
        c->gen_pos_state = 2;
    }
    c->lights_external = false;
}

// light generation on stream `st` (shaders/light_gen.comp) with parameters `u` into (lights, counter), timed into `t`
int enqueue_light_gen(vr_ctx* c, cudaStream_t st, vr_ctx::TimingSet* t, const vr_ubo& u, vr_algo algo, uint32_t gx, uint32_t gy,
                      uint32_t gz, vr_lightgen_variant variant, vr_ray_light* lights, VrLightCounter* counter) {
    VrPassRange range_fn("vr:light_gen");
    const uint64_t nphot = (uint64_t)gx * 4 * gy * 4 * gz;
    cudaEventRecord(t->ev[0], st);
    t->lg_timed = true;
    if (algo != VR_ALGO_PATH) {  // path_light_gen.comp has an empty main (shaders/path_light_gen.comp:9-11)
        const uint32_t nx = gx * 4, ny = gy * 4, nz = gz;  // local_size 4x4x1 (shaders/light_gen.comp:7)
        VR_CUDA(c, c->photon_counts.reserve(nphot));
        VR_CUDA(c, c->photon_offsets.reserve(nphot));
        VR_CUDA(c, c->lg_scan_tmp.reserve((nphot + 4095) / 4096 + 1));
        VR_CUDA(c, c->lg_slab.reserve(nphot * vrk_lightgen_local_capacity() * 2));
        const int keep = (variant == VR_LG_KEEP_ORIGIN) ? 1 : 0;
        vrk_lightgen(st, c->gdev, u, keep, nx, ny, nz, c->photon_counts.p, c->photon_offsets.p, c->photon_total,
                     c->lg_scan_tmp.p, c->lg_slab.p, lights, counter);
    }
    cudaEventRecord(t->ev[1], st);
    return VR_OK;
}

bool same_frame_inputs(const vr_ubo& a, const vr_ubo& b) {  // everything but frameCount
    vr_ubo x = a, y = b;
    x.frameCount = y.frameCount = 0;
    return memcmp(&x, &y, sizeof(vr_ubo)) == 0;
}

// per-frame scalars -> pinned host memory, without blocking (vr_get_counters waits for ev_scalars)
int enqueue_scalars(vr_ctx* c, cudaStream_t st, bool have_march) {
    vr_ctx::Scalars* h = c->scalars_host;
    VR_CUDA(c, cudaMemcpyAsync(&h->nsamples, c->totals, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    if (have_march) VR_CUDA(c, cudaMemcpyAsync(&h->S_ref, c->wc, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    VR_CUDA(c, cudaMemcpyAsync(&h->light_count, &c->counter->lightCount, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    VR_CUDA(c, cudaMemcpyAsync(&h->overflow, c->expand_overflow, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    VR_CUDA(c, cudaEventRecord(c->ev_scalars, st));
    c->counters_pending = true;
    c->counters_have_march = have_march;
    return VR_OK;
}
int finalize_counters(vr_ctx* c) {
    if (!c->counters_pending) return VR_OK;
    VR_CUDA(c, cudaEventSynchronize(c->ev_scalars));
    c->counters_pending = false;
    const vr_ctx::Scalars& h = *c->scalars_host;
    if (c->counters_have_march) {
        c->last_counters.S_ref = c->cached_S[0] = h.S_ref;
        c->last_counters.S_eff = c->cached_S[1] = h.S_eff;
        c->cached_S_valid = true;
    }
    if (h.overflow) return fail(c, VR_E_STATE, "internal: the expanded light list overflowed its buffer");
    if (c->last_algo == VR_ALGO_PATH) {
        c->last_counters.E = h.S_eff;
        return VR_OK;
    }
    if (c->counters_have_march && h.nsamples != c->nsamples_used)
        return fail(c, VR_E_STATE, "internal: camera sample count %llu differs from the cached %llu", (unsigned long long)h.nsamples,
                    (unsigned long long)c->nsamples_used);
    uint32_t lc = h.light_count;
    if (lc > c->light_capacity) lc = c->light_capacity;
    c->last_counters.light_count = lc;
    c->last_counters.eval_lights = h.neval;
    c->last_counters.E = c->cached_S[1] * h.neval;  // (cached_S: the counters of the march these samples came from)
    return VR_OK;
}

}  // namespace

#ifdef VR_CHECKED
void vrk_check_selftest(cudaStream_t st);
#endif
extern "C" {

// bump the kernel revision (kN) whenever an evaluation kernel changes: profiles/eval_traffic.json entries are stamped with it
#ifdef VR_CHECKED
const char* vr_version(void) { return "libvr 0.2-k2 (sm_100a) VR_CHECKED"; }
// diagnostic build only (not in include/vr.h): violates a check on purpose; the next vr_sync must report mask 0x80000000
int vr_checked_selftest(vr_ctx* c) {
    if (!c) return VR_E_ARG;
    cudaSetDevice(c->device);
    vrk_check_selftest(c->stream);
    return VR_OK;
}
#else
const char* vr_version(void) { return "libvr 0.2-k2 (sm_100a)"; }
#endif

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
    {
        // one step above the lowest priority: the side streams of the evaluation pass (lowest) only get the SM slots that the
        // main stream's kernels leave free
        int lo = 0, hi = 0;
        VR_CREATE_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));  // lo = numerically largest = lowest priority
        VR_CREATE_CUDA(cudaStreamCreateWithPriority(&c->own_stream, cudaStreamNonBlocking, lo - 1 >= hi ? lo - 1 : lo));
    }
    c->stream = c->own_stream;
    c->launch_base = vrk_launch_count();
    VR_CREATE_CUDA(cudaMalloc(&c->counter, sizeof(VrLightCounter)));
    VR_CREATE_CUDA(cudaMemsetAsync(c->counter, 0, sizeof(VrLightCounter), c->stream));
    VR_CREATE_CUDA(cudaMalloc(&c->counter_alt, sizeof(VrLightCounter)));
    VR_CREATE_CUDA(cudaMemsetAsync(c->counter_alt, 0, sizeof(VrLightCounter), c->stream));
    VR_CREATE_CUDA(cudaMalloc(&c->photon_total, sizeof(unsigned long long)));
    VR_CREATE_CUDA(cudaMalloc(&c->totals, 2 * sizeof(unsigned long long)));
    VR_CREATE_CUDA(cudaMallocHost(&c->totals_host, 4 * sizeof(unsigned long long)));
    VR_CREATE_CUDA(cudaMalloc(&c->wc, sizeof(VrWorkCounters)));
    VR_CREATE_CUDA(cudaMalloc(&c->expand_overflow, sizeof(uint32_t)));
    VR_CREATE_CUDA(cudaMemsetAsync(c->expand_overflow, 0, sizeof(uint32_t), c->stream));
    VR_CREATE_CUDA(cudaMemsetAsync(c->totals, 0, 2 * sizeof(unsigned long long), c->stream));
    VR_CREATE_CUDA(cudaMemsetAsync(c->wc, 0, sizeof(VrWorkCounters), c->stream));
    VR_CREATE_CUDA(cudaMallocHost(&c->scalars_host, sizeof(vr_ctx::Scalars)));
    memset(c->scalars_host, 0, sizeof(vr_ctx::Scalars));
    VR_CREATE_CUDA(cudaEventCreateWithFlags(&c->ev_scalars, cudaEventDisableTiming));
    {
        // highest priority: the few light-generation CTAs take the first slots the evaluation kernel frees
        int prio_lo = 0, prio_hi = 0;
        VR_CREATE_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
        VR_CREATE_CUDA(cudaStreamCreateWithPriority(&c->aux_stream, cudaStreamNonBlocking, prio_hi));
    }
    VR_CREATE_CUDA(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    {
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);  // lo = the numerically largest = lowest priority
        VR_CREATE_CUDA(cudaStreamCreateWithPriority(&c->seg_stream, cudaStreamNonBlocking, lo));
        VR_CREATE_CUDA(cudaStreamCreateWithPriority(&c->seg_stream2, cudaStreamNonBlocking, lo));
    }
    VR_CREATE_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    VR_CREATE_CUDA(cudaStreamCreateWithFlags(&c->upload_stream, cudaStreamNonBlocking));
    VR_CREATE_CUDA(cudaEventCreateWithFlags(&c->ev_grid_alt_free, cudaEventDisableTiming));
    VR_CREATE_CUDA(cudaEventCreateWithFlags(&c->ev_grid_alt_free_aux, cudaEventDisableTiming));
    VR_CREATE_CUDA(cudaEventCreateWithFlags(&c->ev_copy_ready, cudaEventDisableTiming));
    VR_CREATE_CUDA(cudaEventCreateWithFlags(&c->ev_copy_done, cudaEventDisableTiming));
    VR_CREATE_CUDA(cudaEventCreateWithFlags(&c->ev_seg_fork, cudaEventDisableTiming));
    VR_CREATE_CUDA(cudaEventCreateWithFlags(&c->ev_seg_join, cudaEventDisableTiming));
    VR_CREATE_CUDA(cudaEventCreateWithFlags(&c->ev_seg_join2, cudaEventDisableTiming));
    VR_CREATE_CUDA(cudaEventCreateWithFlags(&c->ev_lg_done, cudaEventDisableTiming));
    VR_CREATE_CUDA(cudaEventCreateWithFlags(&c->ev_frame_end[0], cudaEventDisableTiming));
    VR_CREATE_CUDA(cudaEventCreateWithFlags(&c->ev_frame_end[1], cudaEventDisableTiming));
    VR_CREATE_CUDA(cudaEventCreateWithFlags(&c->ev_grid_ready, cudaEventDisableTiming));
    VR_CREATE_CUDA(cudaMalloc(&c->eval_redo, sizeof(unsigned long long)));
    VR_CREATE_CUDA(cudaMalloc(&c->seg_counts, 8 * sizeof(unsigned long long)));  // layout: VrSegLists::counts
    VR_CREATE_CUDA(cudaMemsetAsync(c->seg_counts, 0, 8 * sizeof(unsigned long long), c->stream));
    VR_CREATE_CUDA(cudaMemsetAsync(c->eval_redo, 0, sizeof(unsigned long long), c->stream));
    for (int k = 0; k < vr_ctx::kTimingSets; k++)
        for (int i = 0; i < 7; i++) VR_CREATE_CUDA(cudaEventCreate(&c->tset[k].ev[i]));
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
    if (c->aux_stream) cudaStreamSynchronize(c->aux_stream);
    if (c->grid) cudaFree(c->grid);
    c->lg_scan_tmp.release();
    c->lg_slab.release();
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
    c->path_tab.release();
    c->light_steps.release();
    c->light_offsets.release();
    if (c->counter) cudaFree(c->counter);
    if (c->counter_alt) cudaFree(c->counter_alt);
    c->lights_alt.release();
    if (c->photon_total) cudaFree(c->photon_total);
    if (c->totals) cudaFree(c->totals);
    if (c->totals_host) cudaFreeHost(c->totals_host);
    if (c->wc) cudaFree(c->wc);
    if (c->expand_overflow) cudaFree(c->expand_overflow);
    if (c->scalars_host) cudaFreeHost(c->scalars_host);
    if (c->ev_scalars) cudaEventDestroy(c->ev_scalars);
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->copy_stream) {
        cudaStreamSynchronize(c->copy_stream);
        cudaStreamDestroy(c->copy_stream);
    }
    if (c->upload_stream) cudaStreamDestroy(c->upload_stream);
    if (c->ev_grid_alt_free) cudaEventDestroy(c->ev_grid_alt_free);
    if (c->ev_grid_alt_free_aux) cudaEventDestroy(c->ev_grid_alt_free_aux);
    if (c->grid_alt) cudaFree(c->grid_alt);
    if (c->ev_copy_ready) cudaEventDestroy(c->ev_copy_ready);
    if (c->ev_copy_done) cudaEventDestroy(c->ev_copy_done);
    c->accum_snapshot.release();
    if (c->ev_seg_fork) cudaEventDestroy(c->ev_seg_fork);
    if (c->ev_seg_join) cudaEventDestroy(c->ev_seg_join);
    if (c->ev_seg_join2) cudaEventDestroy(c->ev_seg_join2);
    if (c->seg_stream) cudaStreamDestroy(c->seg_stream);
    if (c->seg_stream2) cudaStreamDestroy(c->seg_stream2);
    if (c->ev_lg_done) cudaEventDestroy(c->ev_lg_done);
    if (c->ev_frame_end[0]) cudaEventDestroy(c->ev_frame_end[0]);
    if (c->ev_frame_end[1]) cudaEventDestroy(c->ev_frame_end[1]);
    if (c->ev_grid_ready) cudaEventDestroy(c->ev_grid_ready);
    if (c->aux_stream) cudaStreamDestroy(c->aux_stream);
    if (c->eval_redo) cudaFree(c->eval_redo);
    if (c->seg_counts) cudaFree(c->seg_counts);
    c->seg_first.release();
    c->seg_pix.release();
    c->seg_list.release();
    c->seg_list2.release();
    c->seg_sched.release();
    c->seg_part16.release();
    c->seg_part8.release();
    c->seg_base.release();
    c->seg_flag.release();
    c->seg16_first.release();
    c->seg16_pix.release();
    c->seg16_flag.release();
    c->seg_flags.release();
    if (c->accum) cudaFree(c->accum);
    if (c->rgba8) cudaFree(c->rgba8);
    if (c->own_frame_color) cudaFree(c->own_frame_color);
    for (int k = 0; k < vr_ctx::kTimingSets; k++)
        for (int i = 0; i < 7; i++)
            if (c->tset[k].ev[i]) cudaEventDestroy(c->tset[k].ev[i]);
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

// header checks shared by vr_set_grid / vr_set_grid_device: b holds the first `have` bytes of a grid of `size` bytes
static int parse_grid_header(vr_ctx* c, const uint8_t* b, uint64_t have, uint64_t size, uint64_t* root_out, uint32_t* tiles_out) {
    if (size < 672 + 64 + 64) return fail(c, VR_E_BADGRID, "buffer too small for a NanoVDB grid (%llu bytes)", (unsigned long long)size);
    if (have < 672 + 64) return fail(c, VR_E_ARG, "grid header too short (%llu bytes)", (unsigned long long)have);
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
    if (root + 64 > have) return fail(c, VR_E_ARG, "grid header too short: the root node ends at byte %llu", (unsigned long long)(root + 64));
    const uint32_t tiles = rd<uint32_t>(b, root + 24);
    if (root + 64 + 32ull * tiles > size) return fail(c, VR_E_BADGRID, "root tile table outside the buffer");
    *root_out = root, *tiles_out = tiles;
    return VR_OK;
}

// the buffer the next grid goes into: the one of the two that no enqueued work reads
// The buffer the next grid goes into. Its last readers are the frames and light generations enqueued before the previous
// install; `copy_stream` != nullptr: the stream that will fill the buffer waits for them on the device and the host does
// not wait (vr_set_grid_device*), nullptr: the host waits (vr_set_grid, which blocks for its own copy anyway).
static int grid_alt_prepare(vr_ctx* c, uint64_t size, cudaStream_t copy_stream) {
    if (c->grid_alt && c->grid_alt_cap < size) {
        if (c->grid_alt_busy) {
            VR_CUDA(c, cudaEventSynchronize(c->ev_grid_alt_free));
            VR_CUDA(c, cudaEventSynchronize(c->ev_grid_alt_free_aux));
        }
        cudaFree(c->grid_alt);
        c->grid_alt = nullptr;
        c->grid_alt_cap = 0;
    }
    if (!c->grid_alt) {
        VR_CUDA(c, cudaMalloc(&c->grid_alt, size));
        c->grid_alt_cap = size;
        c->grid_alt_busy = false;
    }
    if (c->grid_alt_busy) {
        if (copy_stream) {
            VR_CUDA(c, cudaStreamWaitEvent(copy_stream, c->ev_grid_alt_free, 0));
            VR_CUDA(c, cudaStreamWaitEvent(copy_stream, c->ev_grid_alt_free_aux, 0));
        } else {
            VR_CUDA(c, cudaEventSynchronize(c->ev_grid_alt_free));
            VR_CUDA(c, cudaEventSynchronize(c->ev_grid_alt_free_aux));
        }
    }
    return VR_OK;
}
// after an install: everything enqueued so far reads the buffer that has just become the spare one
static int grid_alt_mark(vr_ctx* c) {
    VR_CUDA(c, cudaEventRecord(c->ev_grid_alt_free, c->stream));
    VR_CUDA(c, cudaEventRecord(c->ev_grid_alt_free_aux, c->aux_stream));
    return VR_OK;
}

// the filled grid_alt becomes the grid; b = header bytes on the host
static void grid_alt_install(vr_ctx* c, const uint8_t* b, uint64_t size, uint64_t root, uint32_t tiles) {
    uint8_t* t = c->grid;
    c->grid = c->grid_alt, c->grid_alt = t;
    const uint64_t tc = c->grid_cap;
    c->grid_cap = c->grid_alt_cap, c->grid_alt_cap = tc;
    c->grid_alt_busy = c->grid_alt != nullptr;
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
}

int vr_set_grid(vr_ctx* c, const void* bytes, uint64_t size) {
    VrPassRange range_fn("vr:set_grid");
    if (!c) return VR_E_ARG;
    if (!bytes) return fail(c, VR_E_ARG, "vr_set_grid: NULL buffer");
    const uint8_t* b = (const uint8_t*)bytes;
    uint64_t root = 0;
    uint32_t tiles = 0;
    int rc = parse_grid_header(c, b, size, size, &root, &tiles);
    if (rc != VR_OK) return rc;
    VR_CUDA(c, cudaSetDevice(c->device));
    c->spec.valid = false;  // (a prefetched light generation may still read the current grid: grid_alt_mark covers it)
    // Two device buffers: the new grid goes into the one no enqueued work reads, on an upload stream, while the frames
    // already enqueued keep rendering from the other (kernels take the grid pointer by value at launch). The host waits
    // for the copy (the caller may reuse `bytes` on return); the GPU does not wait for anything.
    rc = grid_alt_prepare(c, size, nullptr);
    if (rc != VR_OK) return rc;
    VR_CUDA(c, cudaMemcpyAsync(c->grid_alt, bytes, size, cudaMemcpyHostToDevice, c->upload_stream));
    VR_CUDA(c, cudaStreamSynchronize(c->upload_stream));
    rc = grid_alt_mark(c);
    if (rc != VR_OK) return rc;
    c->have_grid = false;
    grid_alt_install(c, b, size, root, tiles);
    c->grid_ready_pending = false;  // (the host waited for the copy)
    return VR_OK;
}

static int set_grid_device_impl(vr_ctx* c, const void* device_bytes, uint64_t size, const void* header, uint64_t header_size,
                                bool have_ready_stream, cudaStream_t ready_stream) {
    if (!c) return VR_E_ARG;
    if (!device_bytes || !header) return fail(c, VR_E_ARG, "vr_set_grid_device: NULL buffer");
    uint64_t root = 0;
    uint32_t tiles = 0;
    int rc = parse_grid_header(c, (const uint8_t*)header, header_size < size ? header_size : size, size, &root, &tiles);
    if (rc != VR_OK) return rc;
    VR_CUDA(c, cudaSetDevice(c->device));
    c->spec.valid = false;
    // no host wait: the stream that copies waits for the spare buffer's last readers (frames and light generations that
    // were enqueued before the previous install), so a host that runs ahead hands the next volume over while the GPU is
    // still one or two frames behind
    rc = grid_alt_prepare(c, size, have_ready_stream ? c->upload_stream : c->stream);
    if (rc != VR_OK) return rc;
    if (have_ready_stream) {
        // the bytes are produced in the order of the caller's `ready_stream` (e.g. the side stream of a replication): the
        // copy into the free grid buffer runs on the upload stream behind that, NEXT TO the frames in flight; whoever
        // reads the new grid (the render stream, the light-generation stream) waits for ev_grid_ready
        VR_CUDA(c, cudaEventRecord(c->ev_grid_ready, ready_stream));
        VR_CUDA(c, cudaStreamWaitEvent(c->upload_stream, c->ev_grid_ready, 0));
        VR_CUDA(c, cudaMemcpyAsync(c->grid_alt, device_bytes, size, cudaMemcpyDefault, c->upload_stream));
        VR_CUDA(c, cudaEventRecord(c->ev_grid_ready, c->upload_stream));
        VR_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_grid_ready, 0));
    } else {
        // device -> device (or peer -> device over NVLink) in stream order: behind whatever the caller enqueued on this
        // context's stream to produce device_bytes (e.g. a broadcast), in front of the next frame; no host synchronisation
        VR_CUDA(c, cudaMemcpyAsync(c->grid_alt, device_bytes, size, cudaMemcpyDefault, c->stream));
        VR_CUDA(c, cudaEventRecord(c->ev_grid_ready, c->stream));
    }
    rc = grid_alt_mark(c);
    if (rc != VR_OK) return rc;
    c->have_grid = false;
    grid_alt_install(c, (const uint8_t*)header, size, root, tiles);
    c->grid_ready_pending = true;
    return VR_OK;
}
int vr_set_grid_device(vr_ctx* c, const void* device_bytes, uint64_t size, const void* header, uint64_t header_size) {
    return set_grid_device_impl(c, device_bytes, size, header, header_size, false, nullptr);
}
int vr_set_grid_device_from_stream(vr_ctx* c, const void* device_bytes, uint64_t size, const void* header, uint64_t header_size,
                                   void* ready_stream) {
    return set_grid_device_impl(c, device_bytes, size, header, header_size, true, (cudaStream_t)ready_stream);
}

int vr_resize(vr_ctx* c, uint32_t w, uint32_t h) {
    if (!c) return VR_E_ARG;
    if (w == 0 || h == 0 || w > 65536 || h > 65536) return fail(c, VR_E_ARG, "image size %ux%u invalid", w, h);
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    if (c->copy_stream) VR_CUDA(c, cudaStreamSynchronize(c->copy_stream));
    c->copy_pending = false;
    if (c->accum) cudaFree(c->accum), c->accum = nullptr;
    if (c->rgba8) cudaFree(c->rgba8), c->rgba8 = nullptr;
    if (c->own_frame_color) cudaFree(c->own_frame_color), c->own_frame_color = nullptr;
    c->frame_color = nullptr;
    memset(&c->fc_out, 0, sizeof(c->fc_out));
    const uint64_t npix = (uint64_t)w * h;
    VR_CUDA(c, cudaMalloc(&c->accum, npix * sizeof(float4)));
    VR_CUDA(c, cudaMalloc(&c->rgba8, npix * sizeof(uchar4)));
    VR_CUDA(c, cudaMalloc(&c->own_frame_color, npix * sizeof(float)));
    c->frame_color = c->own_frame_color;
    c->fc_out.local = c->frame_color;
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
    VR_CUDA(c, cudaStreamSynchronize(c->aux_stream));
    c->spec.valid = false;
    c->lights.release();
    c->lights_alt.release();
    VR_CUDA(c, c->lights.reserve(n));
    VR_CUDA(c, c->lights_alt.reserve(n));
    VR_CUDA(c, cudaMemsetAsync(c->lights.p, 0, c->lights.cap * sizeof(vr_ray_light), c->stream));
    VR_CUDA(c, cudaMemsetAsync(c->lights_alt.p, 0, c->lights_alt.cap * sizeof(vr_ray_light), c->stream));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));  // light generation runs on another stream and must not race these memsets
    c->light_capacity = n;
    if (c->ubo.maxLights > n) c->ubo.maxLights = n;
    int rc = vr_reset_lights(c);
    if (rc != VR_OK) return rc;
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    return VR_OK;
}

int vr_reset_lights(vr_ctx* c) {
    if (!c) return VR_E_ARG;
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaMemsetAsync(c->counter, 0, sizeof(VrLightCounter), c->stream));
    c->lights_bound = 0;
    c->gen_pos_state = 0;
    return VR_OK;
}

int vr_write_lights(vr_ctx* c, const vr_ray_light* lights, uint32_t count) {
    if (!c) return VR_E_ARG;
    if (count > c->light_capacity) return fail(c, VR_E_CAPACITY, "%u lights exceed the capacity %u", count, c->light_capacity);
    if (count && !lights) return fail(c, VR_E_ARG, "vr_write_lights: NULL lights");
    VR_CUDA(c, cudaSetDevice(c->device));
    if (count) VR_CUDA(c, cudaMemcpyAsync(c->lights.p, lights, (size_t)count * sizeof(vr_ray_light), cudaMemcpyHostToDevice, c->stream));
    c->lights_external = true;
    c->external_count = count;
    c->lights_bound = count;
    c->gen_pos_state = 2;  // arbitrary end points: no bound on the ray/beam segment lengths
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

// `t += step` stops advancing once step < ulp(t)/2, i.e. beyond ~1.6e7 steps per ray: the shader loops `while (t < tmax ...)`
// (light_gen.comp:51, point_compute_color.comp:68) and ours would then never end. A ray that starts at `origin_world`
// never has tmax beyond min(rayMaxDistance, distance to the far corner of the grid's box) (intersectAABB clips it once
// and the light-generation walk never recomputes it, light_gen.comp:38-49), so that distance over the step bounds the
// step count of the camera march, of the photon walk (its t restarts at 0 on every scatter) and of path mode alike.
static int check_step_count(vr_ctx* c, const float origin_world[3], const char* what) {
    const VrGridDev& g = c->gdev;
    if (!(c->ubo.rayMarchingStepSize > 0.0f)) return fail(c, VR_E_ARG, "rayMarchingStepSize must be > 0 (the shader loop never ends at 0, src/main.cpp:684)");
    const double sx = origin_world[0] - g.vec[0], sy = origin_world[1] - g.vec[1], sz = origin_world[2] - g.vec[2];
    double far2 = 0.0;
    for (int a = 0; a < 3; a++) {
        const double ci = sx * g.inv[3 * a] + sy * g.inv[3 * a + 1] + sz * g.inv[3 * a + 2];
        const double lo = fabs(ci - (double)g.bmin[a]), hi = fabs((double)g.bmax[a] + 1.0 - ci);
        far2 += (lo > hi ? lo : hi) * (lo > hi ? lo : hi);
    }
    double reach = sqrt(far2);
    const double maxd = (double)c->ubo.rayMaxDistance;
    if (maxd == maxd && maxd < reach) reach = maxd > 0.0 ? maxd : 0.0;
    if (!(reach / (double)c->ubo.rayMarchingStepSize <= 8.0e6))
        return fail(c, VR_E_ARG, "rayMarchingStepSize %g is too small for this grid and %s: more than 8e6 steps per ray, the FP32 march would stop advancing",
                    (double)c->ubo.rayMarchingStepSize, what);
    return VR_OK;
}

static int check_light_gen_args(vr_ctx* c, vr_algo algo, uint32_t gx, uint32_t gy, uint32_t gz) {
    if (!c->have_grid) return fail(c, VR_E_STATE, "no grid: call vr_set_grid first");
    if ((int)algo < 0 || (int)algo > 4) return fail(c, VR_E_ARG, "unknown algorithm %d", (int)algo);
    if (gx == 0 || gy == 0 || gz == 0) return fail(c, VR_E_ARG, "dispatch dimensions must be >= 1");
    const uint64_t nphot = (uint64_t)gx * 4 * gy * 4 * gz;
    if (nphot > (1ull << 28)) return fail(c, VR_E_ARG, "%llu photons exceed the supported dispatch size", (unsigned long long)nphot);
    if (c->ubo.maxLights > c->light_capacity) return fail(c, VR_E_CAPACITY, "maxLights %u exceeds capacity %u", c->ubo.maxLights, c->light_capacity);
    if (algo != VR_ALGO_PATH) return check_step_count(c, c->ubo.lightSourceWorldPos, "light source");  // path_light_gen.comp walks nothing
    return VR_OK;
}

// largest slot total of a single-pass march for which tables and launches are sized by the slot total instead of a
// blocking read of the sample count (tables: about 10 B per slot)
constexpr uint64_t kBoundSizingMaxSlots = 64ull << 20;

// buffers of the ray-segment evaluation for a march with `nreal` samples (tables, index lists, flags, work distribution)
static int reserve_seg_tables(vr_ctx* c, uint64_t nreal, uint64_t npix) {
    const uint64_t n8 = nreal / 8 + 1;
    VR_CUDA(c, c->seg_first.reserve(n8));
    VR_CUDA(c, c->seg_pix.reserve(n8));
    VR_CUDA(c, c->seg_flag.reserve(n8));
    VR_CUDA(c, c->seg16_first.reserve(nreal / 16 + 1));
    VR_CUDA(c, c->seg16_pix.reserve(nreal / 16 + 1));
    VR_CUDA(c, c->seg16_flag.reserve(nreal / 16 + 1));
    VR_CUDA(c, c->seg_flags.reserve(nreal / 8 + 2));
    VR_CUDA(c, c->seg_list.reserve(nreal + 1));
    VR_CUDA(c, c->seg_list2.reserve(nreal + 1));
    VR_CUDA(c, c->seg_sched.reserve(2 * (nreal / 8 / 128 + 4)));  // per table: one word per chunk (+ spare)
    VR_CUDA(c, c->seg_base.reserve(npix + 1));
    VR_CUDA(c, c->seg_part16.reserve(vrk_seg_part_floats(16)));
    VR_CUDA(c, c->seg_part8.reserve(vrk_seg_part_floats(8)));
    c->seg_reserved_for = nreal;
    c->seg_reserved_npix = npix;
    return VR_OK;
}
static void fill_seg_lists(vr_ctx* c, uint64_t nreal, VrSegLists* out) {
    VrSegLists& sl = *out;
    sl.seg_first = c->seg_first.p, sl.seg_pix = c->seg_pix.p, sl.seg_flag = c->seg_flag.p;
    sl.seg16_first = c->seg16_first.p, sl.seg16_pix = c->seg16_pix.p, sl.seg16_flag = c->seg16_flag.p;
    sl.list = c->seg_list.p, sl.list2 = c->seg_list2.p, sl.counts = c->seg_counts;
    const uint64_t per = nreal / 8 / 128 + 4;  // chunks of at least 128 entries (two per thread), at most nreal / 8 entries
    sl.sched16 = c->seg_sched.p, sl.sched8 = c->seg_sched.p + per, sl.sched_bytes = 2 * per * sizeof(uint32_t);
    sl.part16 = c->seg_part16.p, sl.part8 = c->seg_part8.p;
    sl.seg_base = c->seg_base.p, sl.flags = c->seg_flags.p, sl.flag_all = c->seg_sched.p;  // word 0 of the work counters: zeroed with them at the start of every pass
    sl.list_capacity = (uint32_t)nreal;
    sl.cap16 = (uint32_t)(nreal / 16 + 1), sl.cap8 = (uint32_t)(nreal / 8 + 1), sl.cap_flags = (uint32_t)(nreal / 8 + 2);
    sl.sched_words = (uint32_t)per;
    sl.part16_floats = c->seg_part16.cap, sl.part8_floats = c->seg_part8.cap;
    sl.nslots = c->samples.cap;
}

// everything vr_dispatch_compute_color rejects that does not depend on the tile: checked before anything is enqueued
static int check_gather_args(vr_ctx* c, vr_algo algo) {
    if (!c->have_grid) return fail(c, VR_E_STATE, "no grid: call vr_set_grid first");
    if (!c->accum) return fail(c, VR_E_STATE, "no image: call vr_resize first");
    if ((int)algo < 0 || (int)algo > 4) return fail(c, VR_E_ARG, "unknown algorithm %d", (int)algo);
    if ((algo == VR_ALGO_RAY || algo == VR_ALGO_BEAM) && !(c->ubo.lightRayStepSize > 0.0f))
        return fail(c, VR_E_ARG, "lightRayStepSize must be > 0 for ray/beam (uint(len/0) is undefined, ray_compute_color.comp:15)");
    return check_step_count(c, c->ubo.cameraPos, "camera");
}

int vr_dispatch_light_gen(vr_ctx* c, vr_algo algo, uint32_t gx, uint32_t gy, uint32_t gz, vr_lightgen_variant variant) {
    if (!c) return VR_E_ARG;
    int rc = check_light_gen_args(c, algo, gx, gy, gz);
    if (rc != VR_OK) return rc;
    VR_CUDA(c, cudaSetDevice(c->device));
    vr_ctx::TimingSet* t = nullptr;
    rc = open_timing_set(c, &t);
    if (rc != VR_OK) return rc;
    if (c->spec.valid) {  // a prefetched light generation shares the photon scratch buffers: let it drain, then drop it
        VR_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_lg_done, 0));
        c->spec.valid = false;
    }
    rc = enqueue_light_gen(c, c->stream, t, c->ubo, algo, gx, gy, gz, variant, c->lights.p, c->counter);
    if (rc != VR_OK) return rc;
    note_generation(c, c->ubo, algo);
    return check_stream(c);
}

int vr_dispatch_compute_color(vr_ctx* c, vr_algo algo, const vr_tile* tile_or_null) {
    if (!c) return VR_E_ARG;
    int rc = check_gather_args(c, algo);
    if (rc != VR_OK) return rc;
    vr_tile tile;
    rc = validate_tile(c, tile_or_null, &tile);
    if (rc != VR_OK) return rc;
    VR_CUDA(c, cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    const float tan_scale = vr_fov_scale(c->ubo.fov);
    const uint64_t npix = (uint64_t)tile.w * tile.h;
    vr_ctx::TimingSet* t = nullptr;
    rc = open_timing_set(c, &t);
    if (rc != VR_OK) return rc;
    rc = finalize_counters(c);  // a pending error of the previous frame surfaces here
    if (rc != VR_OK) return rc;
    memset(&c->last_counters, 0, sizeof(c->last_counters));
    c->last_algo = algo;

    VrPassRange range(algo == VR_ALGO_PATH ? "vr:path" : "vr:march");
    if (algo == VR_ALGO_PATH) {
        cudaEventRecord(t->ev[2], st);
        cudaEventRecord(t->ev[3], st);
        if (c->lg_in_flight) {
            VR_CUDA(c, cudaStreamWaitEvent(st, c->ev_lg_done, 0));
            c->lg_in_flight = false;
        }
        cudaEventRecord(t->ev[4], st);
        VR_CUDA(c, cudaMemsetAsync(c->wc, 0, sizeof(VrWorkCounters), st));
        // shadow-term table (see k_path_table): usable when the voxel next to the light is one of its 27 neighbours
        // (step <= 1 index unit) and the step count to the farthest bbox corner is moderate
        uint32_t nmax = 0;
        if (c->ubo.rayMarchingStepSize <= 1.0f) {
            const VrGridDev& g = c->gdev;
            double ls[3], far2 = 0.0;
            for (int a = 0; a < 3; a++) {
                const double sx = c->ubo.lightSourceWorldPos[0] - g.vec[0], sy = c->ubo.lightSourceWorldPos[1] - g.vec[1],
                             sz = c->ubo.lightSourceWorldPos[2] - g.vec[2];
                ls[a] = sx * g.inv[3 * a] + sy * g.inv[3 * a + 1] + sz * g.inv[3 * a + 2];
                const double lo = fabs(ls[a] - (double)g.bmin[a]), hi = fabs((double)g.bmax[a] + 1.0 - ls[a]);
                far2 += (lo > hi ? lo : hi) * (lo > hi ? lo : hi);
            }
            const double n = sqrt(far2) * 1.001 / (double)c->ubo.rayMarchingStepSize + 4.0;
            if (n < 65536.0) nmax = (uint32_t)n;
        }
        if (nmax) VR_CUDA(c, c->path_tab.reserve((uint64_t)27 * (nmax + 1)));
        vrk_path(st, c->gdev, c->ubo, tan_scale, tile, c->fc_out, c->accum, c->rgba8, c->wc, nmax ? c->path_tab.p : nullptr, nmax, c->accumulate_mode);
        cudaEventRecord(t->ev[5], st);
        cudaEventRecord(t->ev[6], st);
    cudaEventRecord(c->ev_frame_end[c->pair_parity], st);  // the last reader of this (lights, counter) pair so far
    c->frame_end_recorded[c->pair_parity] = true;
        t->pending = true, t->open = false;
        rc = enqueue_scalars(c, st, true);
        if (rc != VR_OK) return rc;
        return check_stream(c);
    }

    // ---- camera march first: it does not depend on the lights, so it overlaps light generation on the other stream
    cudaEventRecord(t->ev[2], st);
    MarchKey key;
    key.grid_gen = c->grid_gen, key.tile = tile;
    key.dim[0] = c->width, key.dim[1] = c->height;
    memcpy(key.cam, c->ubo.cameraPos, sizeof(key.cam));
    key.fov = c->ubo.fov, key.maxd = c->ubo.rayMaxDistance, key.step = c->ubo.rayMarchingStepSize, key.absorb = c->ubo.absorptionCoefficient;
    key.trilinear = c->phys.trilinear;
    key.valid = true;
    // Single-pass march (default with the ray-segment evaluation, which addresses samples through tables): slots sized by
    // a lookup-free bound of each pixel's step count, so the count pass of the two-pass march (40 % of it) is not needed.
    // The exact kernels (variants 0-3) and the opt-in physics walk a compact sample array: two passes for them.
    key.gaps = (c->march_single && c->eval_variant >= 8 && !(c->phys.phase_hg || c->phys.light_transmittance)) ? 1u : 0u;
    if (key.gaps && key == c->gaps_refused) key.gaps = 0;
    bool same_key = (key == c->march_key);  // same rays over the same grid as the previous gather
    const bool reuse = c->march_cache && same_key;
    uint64_t nsamples = c->cached_samples, nreal = c->cached_real;
    int built_by_march = -1;  // 0 / 1: the march wrote the segment tables itself (without / with 16-sample pairs)
    if (!same_key) c->cached_S_valid = false;  // (set again when this march's counters arrive)
    if (!reuse) {
        VR_CUDA(c, c->px_counts.reserve(npix));
        VR_CUDA(c, c->px_offsets.reserve(npix));
        VR_CUDA(c, c->scan_tmp.reserve((npix + 4095) / 4096 + 1));
        VR_CUDA(c, cudaMemsetAsync(c->wc, 0, sizeof(VrWorkCounters), st));
        BoundKey bk;
        bk.tile = tile, bk.dim[0] = c->width, bk.dim[1] = c->height;
        memcpy(bk.cam, c->ubo.cameraPos, sizeof(bk.cam));
        bk.fov = c->ubo.fov, bk.maxd = c->ubo.rayMaxDistance, bk.step = c->ubo.rayMarchingStepSize;
        memcpy(bk.bmin, c->gdev.bmin, sizeof(bk.bmin)), memcpy(bk.bmax, c->gdev.bmax, sizeof(bk.bmax));
        memcpy(bk.mat, c->gdev.mat, sizeof(bk.mat)), memcpy(bk.inv, c->gdev.inv, sizeof(bk.inv)), memcpy(bk.vec, c->gdev.vec, sizeof(bk.vec));
        bk.valid = true;
        // the slots depend on the rays and on the grid's box and map only: a new grid of the same shape (an animated
        // volume) keeps them, and the frame needs neither the bound kernel nor the blocking read of the total
        const bool same_slots = key.gaps && (bk == c->bound_key);
        if (!key.gaps) c->bound_key.valid = false;  // the two-pass march overwrites px_offsets
        if (key.gaps && !same_slots) {
            c->march_key.valid = false;
            c->bound_key.valid = false;
            vrk_march_bound(st, c->gdev, c->ubo, tan_scale, tile, c->px_counts.p);
            vrk_scan_u32(st, c->px_counts.p, c->px_offsets.p, c->totals + 0, npix, c->scan_tmp.p);
            VR_CUDA(c, cudaMemcpyAsync(c->totals_host, c->totals, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
            VR_CUDA(c, cudaStreamSynchronize(st));
            rc = check_stream(c);
            if (rc != VR_OK) return rc;
            nsamples = c->totals_host[0];
            if (nsamples >= (1ull << 32) - 4096) {  // sample indices are 32-bit in the segment tables
                c->gaps_refused = key;
                key.gaps = 0;
            } else {
                VR_CUDA(c, c->samples.reserve(nsamples + 1));
                c->bound_key = bk;
                c->bound_extent = nsamples;
            }
        }
        if (key.gaps) nsamples = c->bound_extent;
        if (key.gaps) {
            // The slots (offsets, total) are valid: the march alone runs. What sizes the segment tables and the evaluation
            // launches is an upper bound `nreal` of the sample count:
            //  * same rays over the same grid as the previous gather (the steady state of a progressive render; the
            //    reference marches every frame): the previous frame's count, exact;
            //  * new grid or new rays, small frame (an image band of a multi-GPU run, an animated volume): the slot
            //    total itself -- every slot could hold a sample -- so the frame needs NO host round trip and frames
            //    pipeline; launches carry some empty CTAs, which at this size costs less than the bubble of a blocking read;
            //  * new grid or new rays, large frame: one blocking read of the count after the march (the empty CTAs of a
            //    4x oversized launch would cost about 0.25 ms on a 1080p frame).
            // In the first two cases the table sizes are known before the march, and the march writes the tables itself
            // (one launch and its gap less between the march and the evaluation).
            const int pairs = (algo == VR_ALGO_SPHERE || algo == VR_ALGO_BEAM) ? 0 : 1;
            bool read_count = false;
            if (same_key) {
                if (!c->real_exact && c->cached_S_valid && c->cached_S[1] <= nreal) nreal = c->cached_S[1], c->real_exact = true;
            } else if (c->bound_extent <= kBoundSizingMaxSlots) {
                nreal = c->bound_extent, c->real_exact = false;
            } else {
                read_count = true;
            }
            if (!read_count && nreal > 0 && nreal < 0xffffffffull) {
                if (c->seg_reserved_for < nreal || c->seg_reserved_npix < npix) {
                    rc = reserve_seg_tables(c, nreal, npix);
                    if (rc != VR_OK) return rc;
                }
                VrSegLists sl;
                fill_seg_lists(c, nreal, &sl);
                vrk_march_build(st, c->gdev, c->ubo, tan_scale, tile, (int)c->phys.trilinear, c->px_counts.p, c->px_offsets.p, c->samples.p,
                                c->wc, sl, pairs);
                built_by_march = pairs;
            } else {
                vrk_march(st, c->gdev, c->ubo, tan_scale, tile, 2, (int)c->phys.trilinear, c->px_counts.p, c->px_offsets.p, c->samples.p, c->wc);
            }
            if (read_count) {
                VR_CUDA(c, cudaMemcpyAsync(c->totals_host + 2, &c->wc->S_eff, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
                VR_CUDA(c, cudaStreamSynchronize(st));
                rc = check_stream(c);
                if (rc != VR_OK) return rc;
                nreal = c->totals_host[2];
                c->real_exact = true;
            }
        } else {
            vrk_march(st, c->gdev, c->ubo, tan_scale, tile, 0, (int)c->phys.trilinear, c->px_counts.p, nullptr, nullptr, c->wc);
            vrk_scan_u32(st, c->px_counts.p, c->px_offsets.p, c->totals + 0, npix, c->scan_tmp.p);
            if (!same_key) {
                // first gather with these rays: the sample total sizes the buffers, one blocking read. Later frames with
                // the same camera/grid/tile march again (count and fill) but the total is known: no host round trip.
                c->march_key.valid = false;
                VR_CUDA(c, cudaMemcpyAsync(c->totals_host, c->totals, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
                VR_CUDA(c, cudaStreamSynchronize(st));
                rc = check_stream(c);
                if (rc != VR_OK) return rc;
                nsamples = c->totals_host[0];
                VR_CUDA(c, c->samples.reserve(nsamples + 1));
            }
            nreal = nsamples;
            c->real_exact = true;
            vrk_march(st, c->gdev, c->ubo, tan_scale, tile, 1, (int)c->phys.trilinear, c->px_counts.p, c->px_offsets.p, c->samples.p, nullptr);
        }
        c->cached_samples = nsamples;
        c->cached_real = nreal;
        c->march_key = key;
        c->seg_valid = built_by_march >= 0;
        if (built_by_march >= 0) c->seg_pairs = built_by_march;
    }
    VR_CUDA(c, c->contrib.reserve(nsamples + 1));
    c->nsamples_used = nsamples;
    c->nreal_used = nreal;
    cudaEventRecord(t->ev[3], st);

    // ---- light list for the evaluation kernels (needs the lights: join the light-generation stream)
    range.next("vr:light_list");
    if (c->lg_in_flight) {
        VR_CUDA(c, cudaStreamWaitEvent(st, c->ev_lg_done, 0));
        c->lg_in_flight = false;
    }
    const bool expand = (algo == VR_ALGO_RAY || algo == VR_ALGO_BEAM);
    // lightCount lives on the device: launches are sized by maxLights (the shader's cap), kernels read the counter
    // (by the bound recorded when the list was generated: the shader loops over lightCount whatever maxLights is now)
    uint32_t nl_launch = c->lights_bound;
    if (nl_launch > c->light_capacity) nl_launch = c->light_capacity;
    VR_CUDA(c, cudaMemsetAsync(c->expand_overflow, 0, sizeof(uint32_t), st));
    uint64_t eval_list_cap = nl_launch;  // upper bound of the evaluation list's length (the length itself lives on the device)
    if (expand) {
        VR_CUDA(c, c->light_steps.reserve(nl_launch + 1));
        VR_CUDA(c, c->light_offsets.reserve(nl_launch + 1));
        VR_CUDA(c, c->scan_tmp.reserve((nl_launch + 4095) / 4096 + 1));
        vrk_prepare_count_steps(st, c->lights.p, c->counter, c->ubo.lightRayStepSize, c->light_steps.p, nl_launch);
        vrk_scan_u32(st, c->light_steps.p, c->light_offsets.p, c->totals + 1, nl_launch, c->scan_tmp.p);
        uint64_t cap = expand_bound(c);
        if (cap == 0) {  // no bound (lights written by the caller, or a tiny lightRayStepSize): read the total
            VR_CUDA(c, cudaMemcpyAsync(c->totals_host + 1, c->totals + 1, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
            VR_CUDA(c, cudaStreamSynchronize(st));
            cap = c->totals_host[1];
            if (cap > (1ull << 32)) return fail(c, VR_E_ARG, "expanded light list of %llu entries is too large (lightRayStepSize too small?)", (unsigned long long)cap);
        }
        eval_list_cap = cap;
        VR_CUDA(c, c->eval_lights.reserve(cap + 1));
        VR_CUDA(c, c->eval_lights2.reserve(2 * cap + 2));
        vrk_prepare_expand(st, c->lights.p, c->counter, c->ubo.lightRayStepSize, c->light_offsets.p, c->eval_lights.p,
                           c->eval_lights2.p, nl_launch, cap, c->expand_overflow);
    } else {
        VR_CUDA(c, c->eval_lights.reserve((uint64_t)nl_launch + 1));
        VR_CUDA(c, c->eval_lights2.reserve(2 * (uint64_t)nl_launch + 2));
        vrk_prepare_simple(st, c->lights.p, c->counter, algo == VR_ALGO_SPHERE ? 1 : 0, c->eval_lights.p, c->eval_lights2.p,
                           nl_launch, c->totals + 1);
    }
    cudaEventRecord(t->ev[4], st);
    rc = enqueue_scalars(c, st, !reuse);
    if (rc != VR_OK) return rc;
    if (reuse) c->last_counters.S_ref = c->cached_S[0], c->last_counters.S_eff = c->cached_S[1];
    if (c->prefetch_next) {
        // next frame's light generation (frameCount + 1) into the other (lights, counter) pair, on the auxiliary stream,
        // underneath this frame's evaluation kernel; timed into the next frame's event set
        vr_ctx::TimingSet* nt = &c->tset[(c->tcur + 1) % vr_ctx::kTimingSets];
        rc = finalize_set(c, *nt);
        if (rc != VR_OK) return rc;
        nt->open = true;
        vr_ubo nu = c->ubo;
        nu.frameCount += c->frame_step;
        VrPassRange prefetch("vr:light_gen(prefetch)");
        VR_CUDA(c, cudaEventRecord(c->ev_fork, st));
        VR_CUDA(c, cudaStreamWaitEvent(c->aux_stream, c->ev_fork, 0));
        VR_CUDA(c, cudaMemsetAsync(c->counter_alt, 0, sizeof(VrLightCounter), c->aux_stream));
        rc = enqueue_light_gen(c, c->aux_stream, nt, nu, algo, c->lg_groups[0], c->lg_groups[1], c->lg_groups[2], c->lg_variant,
                               c->lights_alt.p, c->counter_alt);
        if (rc != VR_OK) return rc;
        VR_CUDA(c, cudaEventRecord(c->ev_lg_done, c->aux_stream));
        c->spec.valid = true;
        c->spec.ubo = nu, c->spec.grid_gen = c->grid_gen, c->spec.variant = (int)c->lg_variant;
        memcpy(c->spec.groups, c->lg_groups, sizeof(c->spec.groups));
    }

    // ---- evaluation (hot) + resolve
    range.next("vr:evaluate");
    const int sphere = (algo == VR_ALGO_SPHERE || algo == VR_ALGO_BEAM) ? 1 : 0;
    if (c->phys.phase_hg || c->phys.light_transmittance)  // opt-in physics: its own (slow, every-rounding) kernel
        vrk_eval_phys(st, sphere, c->gdev, c->ubo, c->phys, c->samples.p, nsamples, c->eval_lights.p, c->totals + 1, c->contrib.p);
    else if (c->eval_variant == 0)
        vrk_eval(st, sphere, c->samples.p, nsamples, c->eval_lights.p, c->totals + 1, c->ubo.beamRadius, c->contrib.p);
    else if (c->eval_variant >= 8 && nsamples < 0xffffffffull) {
        VrSegLists sl;
        if (c->seg_pairs != (sphere ? 0 : 1)) c->seg_valid = false;
        if (!c->seg_valid) {
            rc = reserve_seg_tables(c, nreal, npix);
            if (rc != VR_OK) return rc;
        }
        fill_seg_lists(c, nreal, &sl);
        if (!c->seg_valid) {
            vrk_build_segments(st, c->px_counts.p, c->px_offsets.p, npix, sl, sphere ? 0 : 1);
            c->seg_valid = true;
            c->seg_pairs = sphere ? 0 : 1;
        }
        // near/far boundary: the share of flagged segments grows with the list length times rho^2, so long lists (ray and
        // beam sub-points) get a smaller rho, down to 0.1 (the kernel's own residual check keeps 2*err/rho <= 4e-4)
        float rho = c->eval_near;
        if (eval_list_cap > 1000 && rho > 0.1f) {
            rho *= sqrtf(1000.0f / (float)eval_list_cap);
            if (rho < 0.1f) rho = 0.1f;
        }
        vrk_eval_point_seg(st, c->eval_side_stream ? c->seg_stream : nullptr, c->eval_side_stream ? c->seg_stream2 : nullptr, c->ev_seg_fork, c->ev_seg_join, c->ev_seg_join2, c->gdev, c->ubo, tan_scale, tile, c->samples.p, nreal, npix, c->px_counts.p, c->px_offsets.p, sl,
                           c->eval_lights.p, c->eval_lights2.p, eval_list_cap, c->totals + 1, rho, sphere, c->ubo.beamRadius,
                           c->contrib.p, c->eval_redo);
    } else if (sphere)
        vrk_eval_sphere_packed(st, c->samples.p, nsamples, c->eval_lights2.p, c->totals + 1, c->ubo.beamRadius, c->contrib.p);
    else
        vrk_eval_point_packed(st, c->eval_variant >= 8 ? 2 : c->eval_variant, c->samples.p, nsamples, c->eval_lights2.p, c->totals + 1,
                              c->contrib.p, c->eval_redo);
    cudaEventRecord(t->ev[5], st);
    range.next("vr:resolve");
    vrk_resolve(st, c->ubo, tile, c->px_counts.p, c->px_offsets.p, c->contrib.p, c->counter, c->fc_out, c->accum, c->rgba8, c->accumulate_mode);
    cudaEventRecord(t->ev[6], st);
    cudaEventRecord(c->ev_frame_end[c->pair_parity], st);  // the last reader of this (lights, counter) pair so far
    c->frame_end_recorded[c->pair_parity] = true;
    t->pending = true, t->open = false;
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
    VrPassRange range_fn("vr:frame");
    if (!c) return VR_E_ARG;
    if (!c->have_grid) return fail(c, VR_E_STATE, "no grid: call vr_set_grid first");
    if (!c->accum) return fail(c, VR_E_STATE, "no image: call vr_resize first");
    // every parameter check of both dispatches runs before anything is enqueued or frameCount changes: a rejected call
    // leaves the context as it was
    int rc = check_light_gen_args(c, algo, c->lg_groups[0], c->lg_groups[1], c->lg_groups[2]);
    if (rc != VR_OK) return rc;
    rc = check_gather_args(c, algo);
    if (rc != VR_OK) return rc;
    {
        vr_tile probe;
        rc = validate_tile(c, tile_or_null, &probe);
        if (rc != VR_OK) return rc;
    }
    VR_CUDA(c, cudaSetDevice(c->device));
    const uint32_t frame_before = c->ubo.frameCount, first_before = c->frame_first;
    if (c->frame_first) c->ubo.frameCount = c->frame_first, c->frame_first = 0;  // vr_set_frame_step
    else c->ubo.frameCount += c->frame_step;                                    // src/main.cpp:712
    vr_ctx::TimingSet* t = nullptr;
    rc = open_timing_set(c, &t);
    if (rc != VR_OK) return rc;
    const vr_lightgen_variant variant = c->lg_variant;
    const uint32_t* gr = c->lg_groups;
    const bool adopt = c->spec.valid && algo != VR_ALGO_PATH && memcmp(&c->spec.ubo, &c->ubo, sizeof(vr_ubo)) == 0 &&
                       c->spec.grid_gen == c->grid_gen && memcmp(c->spec.groups, gr, sizeof(c->spec.groups)) == 0 &&
                       c->spec.variant == (int)variant;
    if (adopt) {
        // this frame's lights were generated (or are being generated) on the auxiliary stream while the previous frame
        // was evaluated: counter reset (:722-728) and light generation (:811-814) already happened, into the other pair
        std::swap(c->lights, c->lights_alt);
        std::swap(c->counter, c->counter_alt);
        c->lights_bound = 0, c->gen_pos_state = 0;  // the other pair's counter was reset before its generation
        note_generation(c, c->spec.ubo, algo);
        c->lg_in_flight = true;  // the gather joins on ev_lg_done
        c->pair_parity ^= 1u;
    } else {
        // Light generation (:811-814) on the auxiliary stream into the OTHER (lights, counter) pair, which the frame before
        // the previous one used last: it depends on that frame's end and on the grid's bytes, not on the previous frame, so
        // when the host runs ahead (an animated volume: new grid, then the next frame, while the GPU still evaluates the
        // previous one) it runs underneath the previous frame's evaluation kernel instead of in front of this frame's
        // gather. The counter reset (:722-728) is the memset of that pair's counter. The gather joins on ev_lg_done: the
        // reference's compute->compute barrier (:817-829).
        const uint32_t np = c->pair_parity ^ 1u;
        if (c->frame_end_recorded[np]) VR_CUDA(c, cudaStreamWaitEvent(c->aux_stream, c->ev_frame_end[np], 0));
        if (c->grid_ready_pending) VR_CUDA(c, cudaStreamWaitEvent(c->aux_stream, c->ev_grid_ready, 0));
        std::swap(c->lights, c->lights_alt);
        std::swap(c->counter, c->counter_alt);
        c->pair_parity = np;
        c->lights_bound = 0, c->gen_pos_state = 0;
        VR_CUDA(c, cudaMemsetAsync(c->counter, 0, sizeof(VrLightCounter), c->aux_stream));
    }
    c->spec.valid = false;
    if (c->ubo.frameCount == 1) {  // :731
        rc = vr_clear_image(c);
        if (rc != VR_OK) return rc;
    }
    if (!adopt) {
        rc = enqueue_light_gen(c, c->aux_stream, t, c->ubo, algo, gr[0], gr[1], gr[2], variant, c->lights.p, c->counter);
        if (rc != VR_OK) return rc;
        note_generation(c, c->ubo, algo);
        VR_CUDA(c, cudaEventRecord(c->ev_lg_done, c->aux_stream));
        c->lg_in_flight = true;
    }
    // steady state (same inputs as the previous frame): prefetch the next frame's lights once this frame's list is built
    c->prefetch_next = c->lg_prefetch && algo != VR_ALGO_PATH && c->prev_frame_valid && c->prev_frame_grid_gen == c->grid_gen &&
                       same_frame_inputs(c->prev_frame_ubo, c->ubo) && c->prev_frame_ubo.frameCount + c->frame_step == c->ubo.frameCount;
    c->prev_frame_ubo = c->ubo, c->prev_frame_grid_gen = c->grid_gen, c->prev_frame_valid = true;
    rc = vr_dispatch_compute_color(c, algo, tile_or_null);
    c->prefetch_next = false;
    if (rc != VR_OK) {
        // the frame did not happen (a CUDA or allocation failure past the argument checks): frameCount goes back, and the
        // steady-state bookkeeping starts over so that no prefetched light list is adopted for a frame that never ran
        c->ubo.frameCount = frame_before, c->frame_first = first_before;
        c->prev_frame_valid = false;
    }
    return rc;
}

int vr_reset_accumulation(vr_ctx* c) {
    if (!c) return VR_E_ARG;
    c->ubo.frameCount = 0;
    return VR_OK;
}

int vr_sync(vr_ctx* c) {
    if (!c) return VR_E_ARG;
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaStreamSynchronize(c->aux_stream));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    int rc = finalize_counters(c);
    if (rc != VR_OK) return rc;
    if (const unsigned int mask = vrk_check_flags())  // (always 0 in the product build)
        return fail(c, VR_E_STATE, "VR_CHECKED: device bounds check failed, mask 0x%x (bits: VR_CHK_* in vr_kernels.cu)", mask);
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

int vr_read_image_rgba8_tile(vr_ctx* c, const vr_tile* tile_or_null, uint8_t* out, uint64_t out_pitch_bytes) {
    if (!c) return VR_E_ARG;
    if (!c->accum) return fail(c, VR_E_STATE, "no image: call vr_resize first");
    if (!out) return fail(c, VR_E_ARG, "NULL output buffer");
    vr_tile tile;
    int rc = validate_tile(c, tile_or_null, &tile);
    if (rc != VR_OK) return rc;
    if (out_pitch_bytes < (uint64_t)tile.w * 4) return fail(c, VR_E_ARG, "row pitch %llu is below the tile's row size", (unsigned long long)out_pitch_bytes);
    VR_CUDA(c, cudaSetDevice(c->device));
    const uchar4* src = c->rgba8 + (size_t)tile.y0 * c->width + tile.x0;
    VR_CUDA(c, cudaMemcpy2DAsync(out, (size_t)out_pitch_bytes, src, (size_t)c->width * 4, (size_t)tile.w * 4, tile.h, cudaMemcpyDeviceToHost, c->stream));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    return VR_OK;
}

/* Asynchronous read-back of the float accumulator: the image is snapshotted on the render stream (device-to-device), the
 * copy to the host runs on a copy stream underneath the frames enqueued afterwards. */
int vr_read_accum_f32_async(vr_ctx* c, float* out) {
    if (!c) return VR_E_ARG;
    if (!c->accum) return fail(c, VR_E_STATE, "no image: call vr_resize first");
    if (!out) return fail(c, VR_E_ARG, "NULL output buffer");
    VR_CUDA(c, cudaSetDevice(c->device));
    const size_t bytes = (size_t)c->width * c->height * 16;
    if (c->copy_pending) VR_CUDA(c, cudaEventSynchronize(c->ev_copy_done));  // one read-back in flight: the snapshot is reused
    VR_CUDA(c, c->accum_snapshot.reserve(bytes / 4));
    VR_CUDA(c, cudaMemcpyAsync(c->accum_snapshot.p, c->accum, bytes, cudaMemcpyDeviceToDevice, c->stream));
    VR_CUDA(c, cudaEventRecord(c->ev_copy_ready, c->stream));
    VR_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_copy_ready, 0));
    VR_CUDA(c, cudaMemcpyAsync(out, c->accum_snapshot.p, bytes, cudaMemcpyDeviceToHost, c->copy_stream));
    VR_CUDA(c, cudaEventRecord(c->ev_copy_done, c->copy_stream));
    c->copy_pending = true;
    return VR_OK;
}
/* The same for one tile (multi-GPU: every GPU returns the rows it rendered over its own PCIe link). `out` addresses the
 * tile's first pixel in the caller's image, rows `out_pitch_bytes` apart. One read-back in flight per context. */
int vr_read_accum_f32_tile_async(vr_ctx* c, const vr_tile* tile_or_null, float* out, uint64_t out_pitch_bytes) {
    if (!c) return VR_E_ARG;
    if (!c->accum) return fail(c, VR_E_STATE, "no image: call vr_resize first");
    if (!out) return fail(c, VR_E_ARG, "NULL output buffer");
    vr_tile tile;
    int rc = validate_tile(c, tile_or_null, &tile);
    if (rc != VR_OK) return rc;
    if (out_pitch_bytes < (uint64_t)tile.w * 16) return fail(c, VR_E_ARG, "row pitch %llu is below the tile's row size", (unsigned long long)out_pitch_bytes);
    VR_CUDA(c, cudaSetDevice(c->device));
    if (c->copy_pending) VR_CUDA(c, cudaEventSynchronize(c->ev_copy_done));
    const size_t row = (size_t)tile.w * 16;
    VR_CUDA(c, c->accum_snapshot.reserve((uint64_t)tile.w * tile.h * 4));
    const float4* src = c->accum + (size_t)tile.y0 * c->width + tile.x0;
    VR_CUDA(c, cudaMemcpy2DAsync(c->accum_snapshot.p, row, src, (size_t)c->width * 16, row, tile.h, cudaMemcpyDeviceToDevice, c->stream));
    VR_CUDA(c, cudaEventRecord(c->ev_copy_ready, c->stream));
    VR_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_copy_ready, 0));
    VR_CUDA(c, cudaMemcpy2DAsync(out, (size_t)out_pitch_bytes, c->accum_snapshot.p, row, row, tile.h, cudaMemcpyDeviceToHost, c->copy_stream));
    VR_CUDA(c, cudaEventRecord(c->ev_copy_done, c->copy_stream));
    c->copy_pending = true;
    return VR_OK;
}
int vr_read_wait(vr_ctx* c) {
    if (!c) return VR_E_ARG;
    if (c->copy_pending) {
        VR_CUDA(c, cudaSetDevice(c->device));
        VR_CUDA(c, cudaEventSynchronize(c->ev_copy_done));
        c->copy_pending = false;
    }
    return check_stream(c);
}

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
    VR_CUDA(c, cudaSetDevice(c->device));
    int rc = finalize_counters(c);  // waits for the frame's scalars only (ready before the evaluation kernel starts)
    if (rc != VR_OK) return rc;
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

int vr_set_frame_color_peers(vr_ctx* c, void* local, void* const* peers, uint32_t n) {
    if (!c) return VR_E_ARG;
    if (!c->accum) return fail(c, VR_E_STATE, "no image: call vr_resize first");
    if (n > VR_MAX_PEERS) return fail(c, VR_E_ARG, "at most %d peer buffers", VR_MAX_PEERS);
    if (n && !peers) return fail(c, VR_E_ARG, "vr_set_frame_color_peers: NULL peer list");
    // no synchronisation: the switch takes effect for kernels enqueued from now on (callers alternate two buffers)
    c->frame_color = local ? (float*)local : c->own_frame_color;
    memset(&c->fc_out, 0, sizeof(c->fc_out));
    c->fc_out.local = c->frame_color;
    for (uint32_t q = 0; q < n; q++) {
        if (!peers[q]) return fail(c, VR_E_ARG, "vr_set_frame_color_peers: NULL peer pointer");
        if (peers[q] == (void*)c->frame_color) continue;  // this GPU's own entry
        c->fc_out.peer[c->fc_out.npeers++] = (float*)peers[q];
    }
    return VR_OK;
}

int vr_accumulate_frame_color(vr_ctx* c, const vr_tile* tile_or_null) {
    VrPassRange range_fn("vr:accumulate_frame_color");
    if (!c) return VR_E_ARG;
    if (!c->accum) return fail(c, VR_E_STATE, "no image: call vr_resize first");
    if (c->ubo.frameCount == 0) return fail(c, VR_E_STATE, "vr_accumulate_frame_color: no frame rendered yet (frameCount == 0)");
    vr_tile tile;
    int rc = validate_tile(c, tile_or_null, &tile);
    if (rc != VR_OK) return rc;
    VR_CUDA(c, cudaSetDevice(c->device));
    vrk_accumulate(c->stream, c->ubo, tile, c->frame_color, c->accum, c->rgba8, c->accumulate_mode);
    return check_stream(c);
}

int vr_get_timings_sum(vr_ctx* c, vr_timings* sum, uint64_t* frames, int reset) {
    if (!c) return VR_E_ARG;
    VR_CUDA(c, cudaSetDevice(c->device));
    int rc = finalize_timings(c);
    if (rc != VR_OK) return rc;
    if (sum) *sum = c->timings_sum;
    if (frames) *frames = c->timings_frames;
    if (reset) {
        memset(&c->timings_sum, 0, sizeof(c->timings_sum));
        c->timings_frames = 0;
    }
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
    vrk_march(c->stream, c->gdev, c->ubo, vr_fov_scale(c->ubo.fov), tile, 0, (int)c->phys.trilinear, c->px_counts.p, nullptr, nullptr, nullptr);
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
    VR_CUDA(c, cudaStreamSynchronize(c->aux_stream));
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
    if (!c || !result || kind < 0 || kind > 9) return VR_E_ARG;
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
        const double rate = ops / (ms * 1e-3) * (kind == 0 ? 2.0 : (kind == 3 || kind >= 7) ? 4.0 : kind == 5 ? 6.0 : kind == 6 ? 5.0 : 1.0) / 1e12;
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

int vr_set_options(vr_ctx* c, const vr_options* o) {
    if (!c || !o) return VR_E_ARG;
    if (o->phase_hg && !(fabsf(o->hg_g) < 1.0f)) return fail(c, VR_E_ARG, "vr_set_options: |hg_g| must be below 1 (got %g)", (double)o->hg_g);
    c->phys.trilinear = o->trilinear ? 1u : 0u;
    c->phys.phase_hg = o->phase_hg ? 1u : 0u;
    c->phys.hg_g = o->hg_g;
    c->phys.light_transmittance = o->light_transmittance ? 1u : 0u;
    return VR_OK;
}

int vr_get_options(vr_ctx* c, vr_options* o) {
    if (!c || !o) return VR_E_ARG;
    o->trilinear = c->phys.trilinear, o->phase_hg = c->phys.phase_hg, o->hg_g = c->phys.hg_g;
    o->light_transmittance = c->phys.light_transmittance;
    return VR_OK;
}

int vr_trilinear_values(vr_ctx* c, const float* xyz, uint64_t n, float* values_out) {
    if (!c) return VR_E_ARG;
    if (!c->have_grid) return fail(c, VR_E_STATE, "no grid: call vr_set_grid first");
    if (n == 0) return VR_OK;
    if (!xyz || !values_out) return fail(c, VR_E_ARG, "vr_trilinear_values: NULL buffer");
    VR_CUDA(c, cudaSetDevice(c->device));
    float *d_xyz = nullptr, *d_out = nullptr;
    VR_CUDA(c, cudaMalloc(&d_xyz, n * 12));
    cudaError_t e = cudaMalloc(&d_out, n * 4);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_xyz, xyz, n * 12, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) {
        vrk_trilinear(c->stream, c->gdev, d_xyz, n, d_out);
        e = cudaMemcpyAsync(values_out, d_out, n * 4, cudaMemcpyDeviceToHost, c->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d_xyz);
    if (d_out) cudaFree(d_out);
    if (e != cudaSuccess) return fail(c, VR_E_CUDA, "vr_trilinear_values failed: %s", cudaGetErrorString(e));
    return check_stream(c);
}

/* calculateRadius of the reference (src/main.cpp:609-618; its call at :713 is commented out): the progressive
 * photon-mapping radius update for 16 photons per pass, in double precision like the reference */
double vr_progressive_radius(double previous_radius, double alpha) {
    const int photonCount = 4 * 4;
    double squaredRadius = previous_radius * previous_radius;
    for (int k = 1; k < photonCount; ++k) squaredRadius *= static_cast<double>(k + alpha) / k;
    squaredRadius = pow(squaredRadius, 1.0 / photonCount);
    return sqrt(squaredRadius);
}

int vr_set_eval_variant(vr_ctx* c, int variant) {
    if (!c || variant < 0 || (variant > 3 && variant != 8)) return VR_E_ARG;
    c->eval_variant = variant;
    return VR_OK;
}

int vr_set_eval_near_distance(vr_ctx* c, float distance) {
    if (!c || !(distance >= 0.01f) || !(distance < 1.0e6f)) return VR_E_ARG;  // at least the 0.01 of the shader's own guard
    c->eval_near = distance;
    return VR_OK;
}

int vr_get_eval_segment_counts(vr_ctx* c, uint64_t out[3]) {
    if (!c || !out) return VR_E_ARG;
    VR_CUDA(c, cudaSetDevice(c->device));
    unsigned long long h[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    VR_CUDA(c, cudaMemcpyAsync(h, c->seg_counts, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    out[0] = h[5], out[1] = h[1], out[2] = h[2] + h[7];  // h[5] = 2 * (16-sample entries) + (8-sample entries); h[7]: handed over
    return check_stream(c);
}

int vr_read_samples(vr_ctx* c, float* out, uint64_t capacity, uint64_t* count) {
    if (!c) return VR_E_ARG;
    VR_CUDA(c, cudaSetDevice(c->device));
    VR_CUDA(c, cudaStreamSynchronize(c->stream));
    {
        const int frc = finalize_counters(c);  // the sample count may only be known as the frame's S_eff counter
        if (frc != VR_OK) return frc;
    }
    const uint64_t n = (c->real_exact || !c->cached_S_valid) ? c->cached_real : c->cached_S[1];
    if (count) *count = n;
    const uint64_t m = n < capacity ? n : capacity;
    if (m && !out) return fail(c, VR_E_ARG, "vr_read_samples: NULL output buffer");
    if (!m) return VR_OK;
    if (!c->march_key.gaps) {
        VR_CUDA(c, cudaMemcpyAsync(out, c->samples.p, m * sizeof(VrSample), cudaMemcpyDeviceToHost, c->stream));
        VR_CUDA(c, cudaStreamSynchronize(c->stream));
        return VR_OK;
    }
    // single-pass march: the sample buffer has gaps; gather the pixels' runs into a compact copy first
    const vr_tile tile = c->march_key.tile;
    const uint64_t npix = (uint64_t)tile.w * tile.h;
    DevBuf<unsigned long long> compact_offsets, total;
    DevBuf<VrSample> compact;
    int rc = VR_OK;
    do {
        if (compact_offsets.reserve(npix) != cudaSuccess || total.reserve(1) != cudaSuccess || compact.reserve(n + 1) != cudaSuccess ||
            c->scan_tmp.reserve((npix + 4095) / 4096 + 1) != cudaSuccess) {
            rc = fail(c, VR_E_CUDA, "vr_read_samples: out of device memory");
            break;
        }
        vrk_scan_u32(c->stream, c->px_counts.p, compact_offsets.p, total.p, npix, c->scan_tmp.p);
        vrk_compact_samples(c->stream, c->px_counts.p, c->px_offsets.p, compact_offsets.p, npix, c->samples.p, compact.p);
        if (cudaMemcpyAsync(out, compact.p, m * sizeof(VrSample), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
            cudaStreamSynchronize(c->stream) != cudaSuccess)
            rc = fail(c, VR_E_CUDA, "vr_read_samples: copy failed");
    } while (0);
    compact_offsets.release();
    total.release();
    compact.release();
    return rc;
}

int vr_get_march_info(vr_ctx* c, uint64_t out[3]) {
    if (!c || !out) return VR_E_ARG;
    const int rc = finalize_counters(c);  // (waits for the last frame's scalars: the sample count may only be known from them)
    if (rc != VR_OK) return rc;
    out[0] = (c->march_key.valid && c->march_key.gaps) ? 1u : 0u;
    out[1] = c->cached_samples;
    out[2] = (c->real_exact || !c->cached_S_valid) ? c->cached_real : c->cached_S[1];
    return VR_OK;
}

int vr_set_march_single_pass(vr_ctx* c, int enabled) {
    if (!c) return VR_E_ARG;
    c->march_single = enabled != 0;
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

int vr_set_eval_side_stream(vr_ctx* c, int enabled) {
    if (!c) return VR_E_ARG;
    c->eval_side_stream = enabled != 0;
    return VR_OK;
}

int vr_set_frame_step(vr_ctx* c, uint32_t first, uint32_t step) {
    if (!c) return VR_E_ARG;
    if (step == 0 || step > (1u << 20)) return fail(c, VR_E_ARG, "frame step %u out of range", step);
    c->frame_step = step;
    c->frame_first = first;
    c->prev_frame_valid = false;  // the prefetched list (if any) was generated for frameCount + the old step
    return VR_OK;
}

int vr_set_accumulate_mode(vr_ctx* c, int mode) {
    if (!c) return VR_E_ARG;
    if (mode != VR_ACCUM_AVERAGE && mode != VR_ACCUM_SUM) return fail(c, VR_E_ARG, "unknown accumulate mode %d", mode);
    c->accumulate_mode = mode;
    return VR_OK;
}

int vr_finish_sum(vr_ctx* c, uint32_t total_frames) {
    if (!c) return VR_E_ARG;
    if (!c->accum) return fail(c, VR_E_STATE, "no image: call vr_resize first");
    if (total_frames == 0) return fail(c, VR_E_ARG, "vr_finish_sum: total_frames must be >= 1");
    VR_CUDA(c, cudaSetDevice(c->device));
    vrk_finish_sum(c->stream, c->accum, c->rgba8, (uint64_t)c->width * c->height, (float)total_frames);
    c->ubo.frameCount = total_frames;
    return check_stream(c);
}

int vr_set_light_prefetch(vr_ctx* c, int enabled) {
    if (!c) return VR_E_ARG;
    c->lg_prefetch = enabled != 0;
    return VR_OK;
}

int vr_set_march_cache(vr_ctx* c, int enabled) {
    if (!c) return VR_E_ARG;
    c->march_cache = enabled != 0;
    c->march_key.valid = false;
    return VR_OK;
}

int vr_cpu_test_render(vr_ctx* c, uint32_t width, uint32_t height, float* density_out) {
    VrPassRange range_fn("vr:cpu_test");
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
    cudaEventRecord(c->tev[0], c->stream);
    vrk_cpu_test(c->stream, c->gdev, width, height, aspect, scale, d);
    cudaEventRecord(c->tev[1], c->stream);
    cudaError_t e = cudaMemcpyAsync(density_out, d, (size_t)width * height * 4, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    float ms = 0;
    cudaEventElapsedTime(&ms, c->tev[0], c->tev[1]);
    finalize_timings(c);
    memset(&c->timings, 0, sizeof(c->timings));
    c->timings.eval_ms = ms, c->timings.total_ms = ms;
    cudaFree(d);
    if (e != cudaSuccess) return fail(c, VR_E_CUDA, "cpu_test render failed: %s", cudaGetErrorString(e));
    return check_stream(c);
}

}  // extern "C"
