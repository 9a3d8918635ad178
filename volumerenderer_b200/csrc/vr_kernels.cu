// vr_kernels.cu -- hand-written sm_100a kernels for VolumeRenderer's many-light hot path.
//
// Pass structure of one frame (replaces the two Vulkan dispatches of src/main.cpp:811-837):
//
//   light generation   k_lightgen<LOCAL> -> scan -> k_lightgen_compact    shaders/light_gen.comp:9-100
//   light expansion    k_prepare_* (ray/beam sub-points pre-expanded)     shaders/ray_compute_color.comp:13-20
//   camera march       k_march_bound -> scan -> k_march<SINGLE> (or <COUNT> -> scan -> <FILL>)   shaders/point_compute_color.comp:20-94
//   evaluation (HOT)   k_eval<POINT|SPHERE>: samples x lights             shaders/common_functions.h:182-201
//   resolve            k_resolve: ordered per-pixel sum, clamp, average   shaders/point_compute_color.comp:97-105
//
// The reference nests the light loop inside each pixel's ray march. Because the pixel colour is linear in the
// per-sample light sums (C = sum_s w_s * L(p_s), w_s = T*val*step, and T does not depend on the lights), the march
// is split from the light loop: the march emits the samples with val != 0 (skipping val == 0 is exact: the shader
// multiplies the light sum by val), the evaluation kernel is then a perfectly load-balanced all-pairs kernel
// (samples x lights) with the light list staged through shared memory by TMA bulk copies, and the resolve pass adds
// each pixel's terms in march order (the reference's summation order).
//
// Exactness: light generation, the march (sample positions, weights, T cut) and path mode use individually rounded
// FP32 ops in the shader's order -> bit-identical with the CPU checker. The evaluation kernel may contract to FMA
// and uses MUFU.RCP; its results are within a few ulp per term (tolerance 1e-3 relative on the pixel, BASELINE.json).
#include <cuda_runtime.h>
#include <stdint.h>

#include "vr_kernels.cuh"

// ---- bounds checks of the diagnostic build (-DVR_CHECKED). compute-sanitizer cannot be run on this pool, so the indices the
// segment path computes (tables built with atomics, index lists, scratch of the slice units, slots of the sample buffer)
// are checked by the kernels themselves; a violated check sets a bit that vr_sync reports. Compiled out otherwise.
enum {
    VR_CHK_SEG16_TABLE = 0, VR_CHK_SEG8_TABLE, VR_CHK_LIST, VR_CHK_FLAGS, VR_CHK_SLOTS, VR_CHK_SCHED, VR_CHK_PART, VR_CHK_LIGHTS,
    VR_CHK_MARCH_SLOT
};
#ifdef VR_CHECKED
__device__ unsigned int g_vr_check = 0;
#define VR_CHECK(cond, bit)                                  \
    do {                                                     \
        if (!(cond)) atomicOr(&g_vr_check, 1u << (bit));     \
    } while (0)
#else
#define VR_CHECK(cond, bit) \
    do {                    \
    } while (0)
#endif

namespace {

// =====================================================================================================
// light generation
// =====================================================================================================
// One photon per thread; photon index p = ((z*Ny)+y)*Nx + x is the canonical order.
// The shader appends with atomicAdd (light_gen.comp:79-87); a deterministic order needs every photon's light count
// before any light can be placed. Instead of walking every photon twice (count, then emit), the walk runs ONCE:
//   LOCAL    each photon stores its first VR_LG_LOCAL lights in a private slab and counts all its scatter events
//   (scan of the counts -> offsets)
//   k_lightgen_compact   copies slab entries whose global slot (base + offsets[p] + j) is below maxLights
//   OVERFLOW photons with more than VR_LG_LOCAL events (rare: the mean is 2.4) walk again and store directly
// A photon's `from` is the position of its previous STORED light; stored lights are a prefix of the photon's list (the
// slot grows with j), so the slab's from/to chain equals the shader's `previousPosition` bookkeeping.
#define VR_LG_LOCAL 16
enum { VR_LG_MODE_LOCAL = 0, VR_LG_MODE_OVERFLOW = 1 };
template <int MODE>
__global__ void __launch_bounds__(64) k_lightgen(VrGridDev g, vr_ubo u, int keep_origin, uint32_t nx, uint32_t ny,
                                                  uint32_t nz, uint32_t* __restrict__ counts, float4* __restrict__ slab,
                                                  const unsigned long long* __restrict__ offsets,
                                                  vr_ray_light* __restrict__ lights, const VrLightCounter* counter) {
    const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t nphot = (uint64_t)nx * ny * nz;
    if (p >= nphot) return;
    if (MODE == VR_LG_MODE_OVERFLOW && counts[p] <= VR_LG_LOCAL) return;
    const uint32_t x = (uint32_t)(p % nx), y = (uint32_t)((p / nx) % ny), z = (uint32_t)(p / ((uint64_t)nx * ny));
    VrSeed seed = {x * u.frameCount, y * u.frameCount, z * u.frameCount};
    const vr_f3 originW = vr_make3(u.lightSourceWorldPos[0], u.lightSourceWorldPos[1], u.lightSourceWorldPos[2]);
    const float r1 = vr_randf_inc(seed);
    const float r2 = vr_randf_inc(seed);
    const vr_f3 dirW = vr_random_dir(r1, r2);
    vr_f3 o = vr_world_to_index(g, originW);
    vr_f3 d = vr_normalize3(vr_world_to_index_dir(g, dirW));
    float tmin = 0.0f, tmax = u.rayMaxDistance;
    const bool hit = vr_intersect_aabb(o, d, g, tmin, tmax);
    const float step = u.rayMarchingStepSize;
    float T = 1.0f, t = tmin, I = u.photonInitialIntensity;
    vr_f3 cur = vr_add3(o, vr_scale3(d, tmin));
    vr_f3 prev = originW;
    uint32_t n = 0;
    // the shader's atomicAdd continues from the counter's current value (0 after the per-frame reset)
    unsigned long long slot = MODE == VR_LG_MODE_OVERFLOW ? (unsigned long long)counter->lightCount + offsets[p] : 0ull;
    float4* mine = MODE == VR_LG_MODE_LOCAL ? slab + p * (2 * VR_LG_LOCAL) : nullptr;
    VrAcc acc;
    vr_acc_init(acc);
    while (hit && t < tmax && T > 0.001f && I > 0.01f) {
        const int32_t i = (int32_t)vr_floor(cur.x), j = (int32_t)vr_floor(cur.y), k = (int32_t)vr_floor(cur.z);
        if (vr_in_bbox(g, i, j, k)) {
            const float val = vr_lookup(g, acc, i, j, k);
            if (val > 0.0f) {
                const float att = vr_exp(vr_mul(vr_mul(-val, u.absorptionCoefficient), step));
                I = vr_mul(I, att);
                T = vr_mul(T, att);
                const float roll = vr_randf_inc(seed);
                if (roll < u.scatteringProbability) {
                    const float q1 = vr_randf_inc(seed);
                    const float q2 = vr_randf_inc(seed);
                    const vr_f3 nd = vr_random_dir(q1, q2);
                    const vr_f3 w = vr_index_to_world(g, cur);
                    if (MODE == VR_LG_MODE_LOCAL) {
                        if (n < VR_LG_LOCAL) {
                            mine[2 * n] = make_float4(prev.x, prev.y, prev.z, 0.0f);
                            mine[2 * n + 1] = make_float4(w.x, w.y, w.z, I);
                            prev = w;
                        }
                    } else {
                        if (slot < (unsigned long long)u.maxLights) {  // light_gen.comp:80-84
                            float4* dst = reinterpret_cast<float4*>(lights + slot);
                            dst[0] = make_float4(prev.x, prev.y, prev.z, 0.0f);
                            dst[1] = make_float4(w.x, w.y, w.z, I);
                            prev = w;
                        }
                        slot++;
                    }
                    n++;
                    d = nd;
                    if (!keep_origin) {  // light_gen.comp:91-92
                        o = cur;
                        t = 0.0f;
                    }
                }
            }
        }
        t = vr_add(t, step);
        cur = vr_add3(o, vr_scale3(d, t));
    }
    if (MODE == VR_LG_MODE_LOCAL) counts[p] = n;
}
// slab entry j of photon p -> lights[base + offsets[p] + j] while that slot is below maxLights
__global__ void __launch_bounds__(256) k_lightgen_compact(vr_ubo u, uint64_t nphot, const uint32_t* __restrict__ counts,
                                                          const unsigned long long* __restrict__ offsets,
                                                          const float4* __restrict__ slab, vr_ray_light* __restrict__ lights,
                                                          const VrLightCounter* counter) {
    const uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t p = e / VR_LG_LOCAL;
    const uint32_t j = (uint32_t)(e % VR_LG_LOCAL);
    if (p >= nphot || j >= counts[p]) return;
    const unsigned long long slot = (unsigned long long)counter->lightCount + offsets[p] + j;
    if (slot >= (unsigned long long)u.maxLights) return;
    float4* dst = reinterpret_cast<float4*>(lights + slot);
    dst[0] = slab[e * 2];
    dst[1] = slab[e * 2 + 1];
}
// Counter value after the dispatch: every scatter did atomicAdd(lightCount, 1); any index >= maxLights stored
// `lightCount = maxLights` (light_gen.comp:85-87). Net effect: min(old + total, maxLights) once anything was emitted.
__global__ void k_lightgen_finish(vr_ubo u, const unsigned long long* __restrict__ total, VrLightCounter* counter) {
    const unsigned long long t = *total;
    counter->totalEmitted += t;
    if (t == 0) return;
    const unsigned long long after = (unsigned long long)counter->lightCount + t;
    counter->lightCount = (uint32_t)(after < (unsigned long long)u.maxLights ? after : (unsigned long long)u.maxLights);
}

// =====================================================================================================
// exclusive scan u32 -> u64
// =====================================================================================================
__device__ __forceinline__ unsigned long long warp_incl_scan(unsigned long long v) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned long long t = __shfl_up_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) >= o) v += t;
    }
    return v;
}
// block-wide exclusive scan of one value per thread (1024 threads); returns exclusive prefix, total in *block_total
__device__ __forceinline__ unsigned long long block_excl_scan(unsigned long long v, unsigned long long* block_total) {
    __shared__ unsigned long long wsum[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const unsigned long long incl = warp_incl_scan(v);
    if (lane == 31) wsum[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        unsigned long long w = (lane < (int)(blockDim.x >> 5)) ? wsum[lane] : 0ull;
        const unsigned long long wi = warp_incl_scan(w);
        wsum[lane] = wi - w;
        if (lane == 31) *block_total = wi;
    }
    __syncthreads();
    const unsigned long long r = wsum[wid] + incl - v;
    __syncthreads();
    return r;
}
__global__ void __launch_bounds__(1024) k_scan_blocks(const uint32_t* __restrict__ in, unsigned long long* __restrict__ out,
                                                       unsigned long long* __restrict__ tmp, uint64_t n) {
    __shared__ unsigned long long total;
    const uint64_t base = ((uint64_t)blockIdx.x * 1024 + threadIdx.x) * 4;
    uint32_t v[4];
#pragma unroll
    for (int q = 0; q < 4; q++) v[q] = (base + q < n) ? in[base + q] : 0u;
    const unsigned long long s = (unsigned long long)v[0] + v[1] + v[2] + v[3];
    unsigned long long ex = block_excl_scan(s, &total);
#pragma unroll
    for (int q = 0; q < 4; q++) {
        if (base + q < n) out[base + q] = ex;
        ex += v[q];
    }
    if (threadIdx.x == 0) tmp[blockIdx.x] = total;
}
__global__ void __launch_bounds__(1024) k_scan_tops(unsigned long long* __restrict__ tmp, uint64_t nblocks,
                                                     unsigned long long* __restrict__ total_out) {
    __shared__ unsigned long long total;
    unsigned long long carry = 0;
    for (uint64_t b0 = 0; b0 < nblocks; b0 += 1024) {
        const uint64_t idx = b0 + threadIdx.x;
        const unsigned long long v = idx < nblocks ? tmp[idx] : 0ull;
        const unsigned long long ex = block_excl_scan(v, &total);
        if (idx < nblocks) tmp[idx] = carry + ex;
        carry += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total_out = carry;
}
__global__ void __launch_bounds__(1024) k_scan_add(unsigned long long* __restrict__ out, const unsigned long long* __restrict__ tmp,
                                                    uint64_t n) {
    const uint64_t base = ((uint64_t)blockIdx.x * 1024 + threadIdx.x) * 4;
    const unsigned long long add = tmp[blockIdx.x];
#pragma unroll
    for (int q = 0; q < 4; q++)
        if (base + q < n) out[base + q] += add;
}

// =====================================================================================================
// light list expansion
// =====================================================================================================
#define VR_INV_4PI 0.07957747154594767f
// The evaluation kernels read two layouts of the expanded list: VrEvalLight (x,y,z,I/4pi; scalar kernel) and the
// duplicated form (x,x,y,y)(z,z,w,w) whose register pairs feed the FP32x2 instructions directly from broadcast LDS.128.
__device__ __forceinline__ void store_eval_light(VrEvalLight* __restrict__ out, float4* __restrict__ out2, uint64_t i, float x,
                                                 float y, float z, float w) {
    reinterpret_cast<float4*>(out)[i] = make_float4(x, y, z, w);
    out2[2 * i] = make_float4(x, x, y, y);
    out2[2 * i + 1] = make_float4(z, z, w, w);
}
// point: positionTo (point_compute_color.comp:12); sphere: positionFrom (sphere_compute_color.comp:12).
// Launched over the buffer capacity (lightCount lives on the device); *neval receives the list length.
__global__ void k_prepare_simple(const vr_ray_light* __restrict__ lights, const VrLightCounter* __restrict__ counter,
                                 int use_from, VrEvalLight* __restrict__ out, float4* __restrict__ out2, uint32_t capacity,
                                 unsigned long long* __restrict__ neval) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t n = counter->lightCount;
    if (n > capacity) n = capacity;
    if (i == 0) *neval = n;
    if (i >= n) return;
    const float4* src = reinterpret_cast<const float4*>(lights + i);
    const float4 a = src[0], b = src[1];
    const float w = b.w * VR_INV_4PI;
    if (use_from) store_eval_light(out, out2, i, a.x, a.y, a.z, w);
    else store_eval_light(out, out2, i, b.x, b.y, b.z, w);
}
// amountOfSteps = uint(length(to - from) / lightRayStepSize)   (ray_compute_color.comp:13-15)
__device__ __forceinline__ uint32_t ray_steps(const float4 a, const float4 b, float lrs, vr_f3* dir) {
    const vr_f3 rl = vr_make3(vr_sub(b.x, a.x), vr_sub(b.y, a.y), vr_sub(b.z, a.z));
    *dir = vr_normalize3(rl);
    const float q = vr_div(vr_length3(rl), lrs);
    // float -> uint: NaN and negatives give 0; saturate well below the buffer limits
    return (q >= 1.0f) ? (q < 4.0e9f ? (uint32_t)q : 0xFFFFFFFFu) : 0u;
}
__global__ void k_prepare_count_steps(const vr_ray_light* __restrict__ lights, const VrLightCounter* __restrict__ counter,
                                      float lrs, uint32_t* __restrict__ steps, uint32_t capacity) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= capacity) return;
    uint32_t n = counter->lightCount;
    if (n > capacity) n = capacity;
    uint32_t s = 0;
    if (i < n) {
        const float4* src = reinterpret_cast<const float4*>(lights + i);
        vr_f3 dir;
        s = ray_steps(src[0], src[1], lrs, &dir);
    }
    steps[i] = s;
}
// sub-point s of light i: from + (float(s) * lrs) * dir, intensity I / float(n)   (ray_compute_color.comp:16-20)
// One block per light slot (launched over the capacity). *overflow is set when the list does not fit out_capacity.
__global__ void k_prepare_expand(const vr_ray_light* __restrict__ lights, const VrLightCounter* __restrict__ counter,
                                 float lrs, const unsigned long long* __restrict__ offsets, VrEvalLight* __restrict__ out,
                                 float4* __restrict__ out2, uint32_t capacity, uint64_t out_capacity, uint32_t* __restrict__ overflow) {
    const uint32_t i = blockIdx.x;
    uint32_t n = counter->lightCount;
    if (n > capacity) n = capacity;
    if (i >= n) return;
    const float4* src = reinterpret_cast<const float4*>(lights + i);
    const float4 a = src[0], b = src[1];
    vr_f3 dir;
    const uint32_t steps = ray_steps(a, b, lrs, &dir);
    if (steps == 0) return;
    const float pointIntensity = vr_div(b.w, vr_u2f(steps));
    const float i4pi = pointIntensity * VR_INV_4PI;
    const unsigned long long off = offsets[i];
    if (off + steps > out_capacity && threadIdx.x == 0) *overflow = 1u;
    for (uint32_t s = threadIdx.x; s < steps; s += blockDim.x) {
        if (off + s >= out_capacity) break;
        const float f = vr_mul(vr_i2f((int32_t)s), lrs);
        store_eval_light(out, out2, off + s, vr_add(a.x, vr_mul(f, dir.x)), vr_add(a.y, vr_mul(f, dir.y)),
                         vr_add(a.z, vr_mul(f, dir.z)), i4pi);
    }
}

// =====================================================================================================
// camera-ray march
// =====================================================================================================
// 128 threads = 16x8 pixels; each warp is an 8x4 pixel packet (neighbouring rays walk the same leaves, so the
// 4-byte value loads of a warp fall into a handful of 128-byte lines). One thread per pixel.
// TRI: the density at a sample is the trilinear interpolation of the eight surrounding voxel values (opt-in, vr_options);
// a zero tile value then says nothing about the interpolated neighbourhood, so the empty-tile skipping is off.
// An upper bound of a pixel's sample count that needs no lookups: the number of steps its march loop can take, from the
// same FP32 sequence t = max(tmin, 0), t += step while t < tmax (the lookups, the bbox test and the T cut-off only ever
// remove steps). With it the march runs ONCE (MODE 2 below): a scan of the bounds gives every pixel a private, large
// enough slot in the sample buffer, which then has gaps; every later kernel addresses samples through (counts, offsets).
__global__ void __launch_bounds__(128) k_march_bound(VrGridDev g, vr_ubo u, float tan_scale, vr_tile tile, uint32_t* __restrict__ bound) {
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t lx = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const uint32_t ly = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
    if (lx >= tile.w || ly >= tile.h) return;
    const VrCamRay r = vr_camera_ray(g, u, tan_scale, tile.x0 + lx, tile.y0 + ly);
    uint32_t n = 0;
    if (r.hit && r.tmax > 0.0f) {
        const float step = u.rayMarchingStepSize;
        for (float t = vr_max(r.tmin, 0.0f); t < r.tmax; t = vr_add(t, step)) n++;
    }
    bound[(uint64_t)ly * tile.w + lx] = n;
}

// table entries of one warp's 32 pixels (c = the pixel's sample count, off = its first sample, p = tile-local pixel
// index; lanes without a pixel pass c = 0). Called by k_build_segments, and by the single-pass march itself when the
// tables' sizes are known beforehand (same rays over the same grid as the previous frame).
__device__ __forceinline__ void seg_build_warp(const uint32_t c, const uint32_t off, const uint64_t p, const bool has_pixel,
                                               const VrSegLists& sl, const int pairs, const uint32_t lane) {
    // q full segments of 8; r leftovers. pairs: q/2 pairs go to the 16-sample table and an odd last one to the 8-sample
    // table (point-type lists: the FMA-bound kernel gains from sharing the geometry between 16 samples); else all q go
    // to the 8-sample table (sphere-type lists: MUFU-bound, nothing to gain, and the flags stay finer)
    const uint32_t q = c >> 3, r = c & 7u, q16 = pairs ? q >> 1 : 0u, q8 = q - 2u * q16;
    uint32_t v[4] = {q, q16, q8, r};  // inclusive warp scans
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t a = __shfl_up_sync(0xffffffffu, v[i], o);
            if (lane >= (uint32_t)o) v[i] += a;
        }
    }
    unsigned long long base[4] = {0, 0, 0, 0};
    if (lane == 31) {
        unsigned long long* ctr[4] = {&sl.counts[5], &sl.counts[4], &sl.counts[0], &sl.counts[1]};
#pragma unroll
        for (int i = 0; i < 4; i++)
            if (v[i]) base[i] = atomicAdd(ctr[i], (unsigned long long)v[i]);
    }
    const uint32_t mine[4] = {q, q16, q8, r};
#pragma unroll
    for (int i = 0; i < 4; i++) base[i] = __shfl_sync(0xffffffffu, base[i], 31) + (v[i] - mine[i]);
    const uint32_t fb = (uint32_t)base[0];  // the pixel's segments are numbered fb .. fb + q - 1 for the near-light flags
    VR_CHECK(base[1] + q16 <= sl.cap16, VR_CHK_SEG16_TABLE);
    VR_CHECK(base[2] + q8 <= sl.cap8, VR_CHK_SEG8_TABLE);
    VR_CHECK(base[3] + r <= sl.list_capacity, VR_CHK_LIST);
    VR_CHECK((uint64_t)fb + q <= sl.cap_flags, VR_CHK_FLAGS);
    VR_CHECK((uint64_t)off + c <= sl.nslots, VR_CHK_SLOTS);
    if (has_pixel) sl.seg_base[p] = fb;
    for (uint32_t i = 0; i < q16; i++) {
        sl.seg16_first[base[1] + i] = off + 16u * i;
        sl.seg16_pix[base[1] + i] = (uint32_t)p;
        sl.seg16_flag[base[1] + i] = fb + 2u * i;
    }
    for (uint32_t i = 0; i < q8; i++) {
        sl.seg_first[base[2] + i] = off + 16u * q16 + 8u * i;
        sl.seg_pix[base[2] + i] = (uint32_t)p;
        sl.seg_flag[base[2] + i] = fb + 2u * q16 + i;
    }
    for (uint32_t j = 0; j < r; j++) sl.list[base[3] + j] = off + 8u * q + j;
}

// MODE 0: count the pixel's samples (counts, work counters); 1: write them at offsets[pix] (after a scan of the counts:
// compact sample buffer); 2: both in one pass, offsets from a scan of k_march_bound's bounds (sample buffer with gaps)
// BUILD (MODE 2 only): the warp also writes its pixels' entries of the segment tables (what k_build_segments does)
template <int MODE, bool TRI, bool BUILD = false>
__global__ void __launch_bounds__(128) k_march(VrGridDev g, vr_ubo u, float tan_scale, vr_tile tile,
                                                uint32_t* __restrict__ counts, const unsigned long long* __restrict__ offsets,
                                                VrSample* __restrict__ samples, VrWorkCounters* wc, VrSegLists sl, int pairs) {
    constexpr bool FILL = MODE != 0, COUNT = MODE != 1;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t lx = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const uint32_t ly = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
    const bool active = lx < tile.w && ly < tile.h;
    uint32_t n = 0, sref = 0, first_slot = 0;
    if (active) {
        const uint32_t px = tile.x0 + lx, py = tile.y0 + ly;
        const uint64_t pix = (uint64_t)ly * tile.w + lx;
        const VrCamRay r = vr_camera_ray(g, u, tan_scale, px, py);
        if (BUILD) first_slot = (uint32_t)offsets[pix];
        if (r.hit && r.tmax > 0.0f) {
            const float step = u.rayMarchingStepSize;
            float t = vr_max(r.tmin, 0.0f);
            float T = 1.0f;
            VrAcc acc;
            vr_acc_init(acc);
            float4* dst = FILL ? reinterpret_cast<float4*>(samples + offsets[pix]) : nullptr;
            while (t < r.tmax && T > 0.001f) {
                const vr_f3 p = vr_add3(r.o, vr_scale3(r.d, t));
                const int32_t i = (int32_t)vr_floor(p.x), j = (int32_t)vr_floor(p.y), k = (int32_t)vr_floor(p.z);
                if (vr_in_bbox(g, i, j, k)) {
                    int log2dim = 0;
                    const float val = TRI ? vr_trilinear(g, acc, p) : vr_lookup_dim(g, acc, i, j, k, log2dim);
                    sref++;
                    if (val != 0.0f) {
                        if (FILL) {
                            const vr_f3 pw = vr_index_to_world(g, p);
                            dst[n] = make_float4(pw.x, pw.y, pw.z, vr_mul(vr_mul(T, val), step));
                        }
                        n++;
                        T = vr_mul(T, vr_exp(vr_mul(vr_mul(-val, u.absorptionCoefficient), step)));
                    } else if (!TRI && log2dim > 0) {
                        // an empty tile: the following lattice points inside it (and inside the bbox) are in-bbox samples
                        // of density 0 -- counted, nothing else to do. t keeps advancing by single adds (the shader's sequence).
                        const float t_lim = vr_min(vr_tile_exit(g, r.o, r.d, p, t, i, j, k, log2dim), r.tmax);
                        float tn = vr_add(t, step);
                        while (tn < t_lim) {
                            t = tn;
                            sref++;
                            tn = vr_add(t, step);
                        }
                    }
                }
                t = vr_add(t, step);
            }
        }
        if (COUNT) counts[pix] = n;
#ifdef VR_CHECKED
        if (MODE == 2 && FILL && pix + 1 < (uint64_t)tile.w * tile.h) VR_CHECK(n <= offsets[pix + 1] - offsets[pix], VR_CHK_MARCH_SLOT);
#endif
    }
    if (BUILD) seg_build_warp(n, first_slot, (uint64_t)ly * tile.w + lx, active, sl, pairs, lane);
    if (COUNT && wc) {
        unsigned long long a = sref, b = n;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, o);
            b += __shfl_xor_sync(0xffffffffu, b, o);
        }
        if (lane == 0) {
            atomicAdd(&wc->S_ref, a);
            atomicAdd(&wc->S_eff, b);
        }
    }
}

// =====================================================================================================
// evaluation: samples x lights  (the hot kernel)
// =====================================================================================================
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// TMA 1-D bulk copy global -> shared, completion signalled on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float rsqrt_approx(float x) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

constexpr int EVAL_THREADS = 256;
constexpr int EVAL_TILE = 1024;  // lights per shared-memory stage (16 KB)

// Each thread keeps S samples in registers and walks the light list in list order with ONE accumulator per sample
// (the reference's summation order, point_compute_color.comp:11-13). The light list is streamed through a
// two-stage shared-memory ring filled by TMA bulk copies; every LDS.128 of a light is a warp-wide broadcast and is
// amortised over the S samples of the thread.
template <bool SPHERE, int S>
__global__ void __launch_bounds__(EVAL_THREADS) k_eval(const VrSample* __restrict__ samples, uint64_t nsamples,
                                                        const VrEvalLight* __restrict__ lights,
                                                        const unsigned long long* __restrict__ nlights_dev, float radius,
                                                        float* __restrict__ out) {
    __shared__ __align__(128) float4 stage[2][EVAL_TILE];
    const uint64_t nlights = *nlights_dev;  // list length lives on the device (no host round trip per frame)
    __shared__ __align__(8) uint64_t bar[2];
    const uint32_t tid = threadIdx.x;
    const uint64_t base = (uint64_t)blockIdx.x * (EVAL_THREADS * S);
    const uint64_t ntiles = (nlights + EVAL_TILE - 1) / EVAL_TILE;
    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0 && ntiles > 0) {
        const uint32_t cnt = (uint32_t)(nlights < EVAL_TILE ? nlights : EVAL_TILE);
        mbar_expect_tx(&bar[0], cnt * 16u);
        tma_bulk_g2s(&stage[0][0], lights, cnt * 16u, &bar[0]);
    }
    float sx[S], sy[S], sz[S], sw[S], acc[S];
#pragma unroll
    for (int j = 0; j < S; j++) {
        const uint64_t idx = base + (uint64_t)j * EVAL_THREADS + tid;
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        if (idx < nsamples) s = __ldg(reinterpret_cast<const float4*>(samples) + idx);
        sx[j] = s.x, sy[j] = s.y, sz[j] = s.z, sw[j] = s.w, acc[j] = 0.0f;
    }
    for (uint64_t tl = 0; tl < ntiles; tl++) {
        const uint32_t buf = (uint32_t)(tl & 1);
        if (tid == 0 && tl + 1 < ntiles) {  // prefetch the next tile into the other stage (freed by the barrier below)
            const uint64_t first = (tl + 1) * EVAL_TILE;
            const uint32_t cnt = (uint32_t)((nlights - first) < EVAL_TILE ? (nlights - first) : EVAL_TILE);
            mbar_expect_tx(&bar[buf ^ 1], cnt * 16u);
            tma_bulk_g2s(&stage[buf ^ 1][0], lights + first, cnt * 16u, &bar[buf ^ 1]);
        }
        mbar_wait(&bar[buf], (uint32_t)((tl >> 1) & 1));
        const uint64_t first = tl * EVAL_TILE;
        const int cnt = (int)((nlights - first) < EVAL_TILE ? (nlights - first) : EVAL_TILE);
        const float4* __restrict__ L = stage[buf];
#pragma unroll 4
        for (int l = 0; l < cnt; l++) {
            const float4 q = L[l];
#pragma unroll
            for (int j = 0; j < S; j++) {
                if (SPHERE) {
                    // getClosestPointOnSphere + getPointLightContribution (common_functions.h:182-201), same op order:
                    // cur = O + normalize(p - O) * R ; Lv = cur - p
                    const float vx = sx[j] - q.x, vy = sy[j] - q.y, vz = sz[j] - q.z;
                    const float l2 = __fadd_rn(__fadd_rn(__fmul_rn(vx, vx), __fmul_rn(vy, vy)), __fmul_rn(vz, vz));
                    const float inv = rsqrt_approx(l2);
                    const float cx = __fadd_rn(q.x, __fmul_rn(__fmul_rn(vx, inv), radius));
                    const float cy = __fadd_rn(q.y, __fmul_rn(__fmul_rn(vy, inv), radius));
                    const float cz = __fadd_rn(q.z, __fmul_rn(__fmul_rn(vz, inv), radius));
                    const float dx = cx - sx[j], dy = cy - sy[j], dz = cz - sz[j];
                    const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
                    const float r = rcp_approx(d2);
                    acc[j] = (d2 < 0.0001f) ? acc[j] : fmaf(q.w, r, acc[j]);
                } else {
                    const float dx = q.x - sx[j], dy = q.y - sy[j], dz = q.z - sz[j];
                    const float d2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
                    const float r = rcp_approx(d2);
                    acc[j] = (d2 < 0.0001f) ? acc[j] : fmaf(q.w, r, acc[j]);
                }
            }
        }
        __syncthreads();  // everyone is done with stage[buf] before it is refilled two iterations later
    }
#pragma unroll
    for (int j = 0; j < S; j++) {
        const uint64_t idx = base + (uint64_t)j * EVAL_THREADS + tid;
        if (idx < nsamples) out[idx] = sw[j] * acc[j];
    }
}

// ---- packed variant (point lights): two samples per FP32x2 instruction -----------------------------------------
// Blackwell (sm_100) adds packed FP32 instructions (PTX add/mul/fma.rn.f32x2, SASS FADD2/FMUL2/FFMA2) that carry two
// FP32 operations per issue slot. The baseline kernel above is issue-bound (ncu r01: 9.3 instructions per evaluation,
// issue slots 75 % busy, XU pipe 65 %); packing the distance computation of two samples of the same thread halves its
// issue cost (6 -> 3 slots per evaluation), leaving the MUFU.RCP rate (16/clk/SM) as the next ceiling.
// The light list is stored duplicated, (x,x,y,y)(z,z,w,w) (store_eval_light), so that both halves of every register
// pair come straight from two broadcast LDS.128.
constexpr int EVAL2_TILE = 512;  // lights per stage: 512 * 32 B = 16 KB

template <int S>  // S samples per thread, S even
__global__ void __launch_bounds__(EVAL_THREADS) k_eval_point2(const VrSample* __restrict__ samples, uint64_t nsamples,
                                                              const float4* __restrict__ lights2,
                                                              const unsigned long long* __restrict__ nlights_dev,
                                                              float* __restrict__ out) {
    constexpr int P = S / 2;
    const uint64_t nlights = *nlights_dev;
    __shared__ __align__(128) float4 stage[2][2 * EVAL2_TILE];
    __shared__ __align__(8) uint64_t bar[2];
    const uint32_t tid = threadIdx.x;
    const uint64_t base = (uint64_t)blockIdx.x * (EVAL_THREADS * S);
    const uint64_t ntiles = (nlights + EVAL2_TILE - 1) / EVAL2_TILE;
    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0 && ntiles > 0) {
        const uint32_t cnt = (uint32_t)(nlights < EVAL2_TILE ? nlights : EVAL2_TILE);
        mbar_expect_tx(&bar[0], cnt * 32u);
        tma_bulk_g2s(&stage[0][0], lights2, cnt * 32u, &bar[0]);
    }
    // negated sample coordinates (the packed ISA has add, not sub), pairs (2p, 2p+1)
    float2 nsx[P], nsy[P], nsz[P], acc[P];
    float sw[S];
#pragma unroll
    for (int p = 0; p < P; p++) {
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
        const uint64_t i0 = base + (uint64_t)(2 * p) * EVAL_THREADS + tid, i1 = i0 + EVAL_THREADS;
        if (i0 < nsamples) a = __ldg(reinterpret_cast<const float4*>(samples) + i0);
        if (i1 < nsamples) b = __ldg(reinterpret_cast<const float4*>(samples) + i1);
        nsx[p] = make_float2(-a.x, -b.x), nsy[p] = make_float2(-a.y, -b.y), nsz[p] = make_float2(-a.z, -b.z);
        sw[2 * p] = a.w, sw[2 * p + 1] = b.w;
        acc[p] = make_float2(0.f, 0.f);
    }
    for (uint64_t tl = 0; tl < ntiles; tl++) {
        const uint32_t buf = (uint32_t)(tl & 1);
        if (tid == 0 && tl + 1 < ntiles) {
            const uint64_t first = (tl + 1) * EVAL2_TILE;
            const uint32_t cnt = (uint32_t)((nlights - first) < EVAL2_TILE ? (nlights - first) : EVAL2_TILE);
            mbar_expect_tx(&bar[buf ^ 1], cnt * 32u);
            tma_bulk_g2s(&stage[buf ^ 1][0], lights2 + 2 * first, cnt * 32u, &bar[buf ^ 1]);
        }
        mbar_wait(&bar[buf], (uint32_t)((tl >> 1) & 1));
        const uint64_t first = tl * EVAL2_TILE;
        const int cnt = (int)((nlights - first) < EVAL2_TILE ? (nlights - first) : EVAL2_TILE);
        const float4* __restrict__ L = stage[buf];
#pragma unroll 2
        for (int l = 0; l < cnt; l++) {
            const float4 qa = L[2 * l], qb = L[2 * l + 1];
            const float2 X = make_float2(qa.x, qa.y), Y = make_float2(qa.z, qa.w), Z = make_float2(qb.x, qb.y);
            const float wI = qb.z;
#pragma unroll
            for (int p = 0; p < P; p++) {
                const float2 dx = __fadd2_rn(X, nsx[p]);
                const float2 dy = __fadd2_rn(Y, nsy[p]);
                const float2 dz = __fadd2_rn(Z, nsz[p]);
                float2 d2 = __fmul2_rn(dx, dx);
                d2 = __ffma2_rn(dy, dy, d2);
                d2 = __ffma2_rn(dz, dz, d2);
                const float r0 = rcp_approx(d2.x), r1 = rcp_approx(d2.y);
                acc[p].x = (d2.x < 0.0001f) ? acc[p].x : fmaf(wI, r0, acc[p].x);
                acc[p].y = (d2.y < 0.0001f) ? acc[p].y : fmaf(wI, r1, acc[p].y);
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int p = 0; p < P; p++) {
        const uint64_t i0 = base + (uint64_t)(2 * p) * EVAL_THREADS + tid, i1 = i0 + EVAL_THREADS;
        if (i0 < nsamples) out[i0] = sw[2 * p] * acc[p].x;
        if (i1 < nsamples) out[i1] = sw[2 * p + 1] * acc[p].y;
    }
}

// ---- packed variant (sphere lights, also the sub-points of beams) ---------------------------------------------------
// getClosestPointOnSphere + getPointLightContribution (common_functions.h:182-201) on two samples per FP32x2
// instruction. The scalar kernel k_eval<SPHERE> spends ~30 issue slots per evaluation (it keeps every rounding of the
// shader separate); here the roundings that cannot matter are merged:
//   l2  = |p - O|^2 as fmul + 2 fma                      (1 ulp of l2 -> 1 ulp of inv -> R * 6e-8 = 6e-9 in `cur`)
//   s   = inv * R once per evaluation, cur = fma(v, s, O) (inner roundings are at the 1e-8 scale)
//   d2  = |cur - p|^2 as fmul + 2 fma
// What is kept is the ONE rounding that dominates the result: `cur` rounded at the magnitude of the light position
// (ulp(50) = 4e-6, amplified by 1/|cur - p| near the sphere surface), so the kernel tracks the reference's own
// arithmetic noise instead of the ideal (|v| - R)^2. 16 packed + 4 MUFU + 4 guard/accumulate slots per pair of
// evaluations: FMA pipe 17 cycles, XU pipe 16 cycles per evaluation.
// LIST: the kernel evaluates samples[list[i]], i < *nlist_dev (the leftovers and near-light segments of the ray-segment
// kernel); the launch is sized by an upper bound of the list length and CTAs beyond it leave at once.
// STRIDE (LIST only): the CTAs of a fixed small grid walk the list in strides (the hand-over list, normally empty)
template <int S, bool LIST, bool STRIDE = false>
__global__ void __launch_bounds__(EVAL_THREADS, 3) k_eval_sphere2(const VrSample* __restrict__ samples, uint64_t nsamples,
                                                                   const uint32_t* __restrict__ list,
                                                                   const unsigned long long* __restrict__ nlist_begin_dev,
                                                                   const unsigned long long* __restrict__ nlist_dev,
                                                                   const float4* __restrict__ lights2,
                                                                   const unsigned long long* __restrict__ nlights_dev, float radius,
                                                                   float* __restrict__ out) {
    constexpr int P = S / 2;
    if (LIST) {  // entries [*nlist_begin_dev (0 if NULL), *nlist_dev) of the list
        const uint64_t lbeg = nlist_begin_dev ? *nlist_begin_dev : 0;
        list += lbeg;
        nsamples = *nlist_dev - lbeg;
        if ((uint64_t)blockIdx.x * (EVAL_THREADS * S) >= nsamples) return;
    }
    const uint64_t nlights = *nlights_dev;
    __shared__ __align__(128) float4 stage[2][2 * EVAL2_TILE];
    __shared__ __align__(8) uint64_t bar[2];
    const uint32_t tid = threadIdx.x;
    const uint64_t ntiles = (nlights + EVAL2_TILE - 1) / EVAL2_TILE;
    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t it = 0;  // tiles consumed so far by this CTA: stage = it & 1, mbarrier phase parity = (it >> 1) & 1
    uint64_t base = (uint64_t)blockIdx.x * (EVAL_THREADS * S);
    do {
    if (tid == 0 && ntiles > 0) {
        const uint32_t cnt = (uint32_t)(nlights < EVAL2_TILE ? nlights : EVAL2_TILE);
        mbar_expect_tx(&bar[it & 1], cnt * 32u);
        tma_bulk_g2s(&stage[it & 1][0], lights2, cnt * 32u, &bar[it & 1]);
    }
    float2 nsx[P], nsy[P], nsz[P], acc[P];
    float sw[S];
#pragma unroll
    for (int p = 0; p < P; p++) {
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
        const uint64_t i0 = base + (uint64_t)(2 * p) * EVAL_THREADS + tid, i1 = i0 + EVAL_THREADS;
        if (i0 < nsamples) a = __ldg(reinterpret_cast<const float4*>(samples) + (LIST ? (uint64_t)list[i0] : i0));
        if (i1 < nsamples) b = __ldg(reinterpret_cast<const float4*>(samples) + (LIST ? (uint64_t)list[i1] : i1));
        nsx[p] = make_float2(-a.x, -b.x), nsy[p] = make_float2(-a.y, -b.y), nsz[p] = make_float2(-a.z, -b.z);
        sw[2 * p] = a.w, sw[2 * p + 1] = b.w;
        acc[p] = make_float2(0.f, 0.f);
    }
    const float2 nR = make_float2(-radius, -radius);
    for (uint64_t tl = 0; tl < ntiles; tl++, it++) {
        const uint32_t buf = it & 1;
        if (tid == 0 && tl + 1 < ntiles) {
            const uint64_t first = (tl + 1) * EVAL2_TILE;
            const uint32_t cnt = (uint32_t)((nlights - first) < EVAL2_TILE ? (nlights - first) : EVAL2_TILE);
            mbar_expect_tx(&bar[buf ^ 1], cnt * 32u);
            tma_bulk_g2s(&stage[buf ^ 1][0], lights2 + 2 * first, cnt * 32u, &bar[buf ^ 1]);
        }
        mbar_wait(&bar[buf], (it >> 1) & 1);
        const uint64_t first = tl * EVAL2_TILE;
        const int cnt = (int)((nlights - first) < EVAL2_TILE ? (nlights - first) : EVAL2_TILE);
        const float4* __restrict__ L = stage[buf];
#pragma unroll 2
        for (int l = 0; l < cnt; l++) {
            const float4 qa = L[2 * l], qb = L[2 * l + 1];
            const float2 X = make_float2(qa.x, qa.y), Y = make_float2(qa.z, qa.w), Z = make_float2(qb.x, qb.y);
            const float wI = qb.z;
#pragma unroll
            for (int p = 0; p < P; p++) {
                // v = O - p (the negated p - O: squares and the product with the negated radius absorb the sign exactly)
                const float2 vx = __fadd2_rn(X, nsx[p]), vy = __fadd2_rn(Y, nsy[p]), vz = __fadd2_rn(Z, nsz[p]);
                float2 l2 = __fmul2_rn(vx, vx);
                l2 = __ffma2_rn(vy, vy, l2);
                l2 = __ffma2_rn(vz, vz, l2);
                const float2 sc = __fmul2_rn(make_float2(rsqrt_approx(l2.x), rsqrt_approx(l2.y)), nR);  // -R / |p - O|
                // cur = O + normalize(p - O) * R ; Lv = cur - p
                const float2 lx = __fadd2_rn(__ffma2_rn(vx, sc, X), nsx[p]);
                const float2 ly = __fadd2_rn(__ffma2_rn(vy, sc, Y), nsy[p]);
                const float2 lz = __fadd2_rn(__ffma2_rn(vz, sc, Z), nsz[p]);
                float2 d2 = __fmul2_rn(lx, lx);
                d2 = __ffma2_rn(ly, ly, d2);
                d2 = __ffma2_rn(lz, lz, d2);
                const float r0 = rcp_approx(d2.x), r1 = rcp_approx(d2.y);
                acc[p].x = (d2.x < 0.0001f) ? acc[p].x : fmaf(wI, r0, acc[p].x);
                acc[p].y = (d2.y < 0.0001f) ? acc[p].y : fmaf(wI, r1, acc[p].y);
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int p = 0; p < P; p++) {
        const uint64_t i0 = base + (uint64_t)(2 * p) * EVAL_THREADS + tid, i1 = i0 + EVAL_THREADS;
        if (i0 < nsamples) out[LIST ? (uint64_t)list[i0] : i0] = sw[2 * p] * acc[p].x;
        if (i1 < nsamples) out[LIST ? (uint64_t)list[i1] : i1] = sw[2 * p + 1] * acc[p].y;
    }
    base += (uint64_t)gridDim.x * (EVAL_THREADS * S);
    } while (STRIDE && base < nsamples);
}

// ---- speculative variant: no per-evaluation guard in the main pass -------------------------------------------------
// ncu (profiles/r01b): k_eval_point2 is bound by the XU pipe (MUFU.RCP, 82.6 % busy) with the FMA pipe at 72 % and the
// issue slots at 70 %: three resources within 15 % of each other, so the scheduler cannot keep the XU pipe full. The
// guard `d2 < 1e-4 ? 0 : ...` (common_functions.h:190) costs two of the 6.7 issue slots per evaluation (FSETP + a
// predicated scalar FFMA instead of half an FFMA2), yet it fires in about 6 of 32 000 thread blocks per C2 frame (a sample
// within 0.01 world units of a light). This kernel therefore runs the light loop without the guard, tracking the
// minimum d2 each thread has seen with one 3-input FMNMX per two evaluations (sm_100 `min.f32 d,a,b,c`), and
// accumulates with FFMA2. If any thread of the CTA saw d2 < 1e-4 the CTA repeats its light loop with the guarded body
// (the result of the first pass is discarded), so the output is identical to k_eval_point2's in every case.
// Measured on C2 (profiles/README.md, r01c): 16.55 ms vs 17.12 ms. Moving a fraction of the reciprocals to the FMA pipe
// (integer seed + Newton steps) was tried and is slower (17.7-18.5 ms): the FMA pipe is as loaded as the XU pipe.
__device__ __forceinline__ float fmin3(float a, float b, float c) {
    float r;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
template <int P, bool GUARD>
__device__ __forceinline__ void eval_light_packed(const float4 qa, const float4 qb, const float2 (&nsx)[P], const float2 (&nsy)[P],
                                                  const float2 (&nsz)[P], float2 (&acc)[P], float& dmin) {
    const float2 X = make_float2(qa.x, qa.y), Y = make_float2(qa.z, qa.w), Z = make_float2(qb.x, qb.y);
    const float2 W = make_float2(qb.z, qb.w);
#pragma unroll
    for (int p = 0; p < P; p++) {
        const float2 dx = __fadd2_rn(X, nsx[p]);
        const float2 dy = __fadd2_rn(Y, nsy[p]);
        const float2 dz = __fadd2_rn(Z, nsz[p]);
        float2 d2 = __fmul2_rn(dx, dx);
        d2 = __ffma2_rn(dy, dy, d2);
        d2 = __ffma2_rn(dz, dz, d2);
        if (GUARD) {
            const float r0 = rcp_approx(d2.x), r1 = rcp_approx(d2.y);
            acc[p].x = (d2.x < 0.0001f) ? acc[p].x : fmaf(W.x, r0, acc[p].x);
            acc[p].y = (d2.y < 0.0001f) ? acc[p].y : fmaf(W.y, r1, acc[p].y);
        } else {
            dmin = fmin3(dmin, d2.x, d2.y);
            acc[p] = __ffma2_rn(W, make_float2(rcp_approx(d2.x), rcp_approx(d2.y)), acc[p]);
        }
    }
}

// one pass over the whole light list through the two-stage TMA ring; `it` counts the tiles consumed so far by this CTA
// (stage = it & 1, mbarrier phase parity = (it >> 1) & 1) so that a second pass continues the ring where the first ended
template <int P, bool GUARD, int UNROLL>
__device__ __forceinline__ void eval_pass_packed(float4 (&stage)[2][2 * EVAL2_TILE], uint64_t (&bar)[2], const float4* __restrict__ lights2,
                                                 uint64_t nlights, uint32_t& it, const float2 (&nsx)[P], const float2 (&nsy)[P],
                                                 const float2 (&nsz)[P], float2 (&acc)[P], float& dmin) {
    const uint32_t tid = threadIdx.x;
    const uint64_t ntiles = (nlights + EVAL2_TILE - 1) / EVAL2_TILE;
    if (tid == 0 && ntiles > 0) {
        const uint32_t cnt = (uint32_t)(nlights < EVAL2_TILE ? nlights : EVAL2_TILE);
        mbar_expect_tx(&bar[it & 1], cnt * 32u);
        tma_bulk_g2s(&stage[it & 1][0], lights2, cnt * 32u, &bar[it & 1]);
    }
    for (uint64_t tl = 0; tl < ntiles; tl++, it++) {
        const uint32_t buf = it & 1;
        if (tid == 0 && tl + 1 < ntiles) {
            const uint64_t first = (tl + 1) * EVAL2_TILE;
            const uint32_t cnt = (uint32_t)((nlights - first) < EVAL2_TILE ? (nlights - first) : EVAL2_TILE);
            mbar_expect_tx(&bar[buf ^ 1], cnt * 32u);
            tma_bulk_g2s(&stage[buf ^ 1][0], lights2 + 2 * first, cnt * 32u, &bar[buf ^ 1]);
        }
        mbar_wait(&bar[buf], (it >> 1) & 1);
        const uint64_t first = tl * EVAL2_TILE;
        const int cnt = (int)((nlights - first) < EVAL2_TILE ? (nlights - first) : EVAL2_TILE);
        const float4* __restrict__ L = stage[buf];
#pragma unroll UNROLL
        for (int l = 0; l < cnt; l++) eval_light_packed<P, GUARD>(L[2 * l], L[2 * l + 1], nsx, nsy, nsz, acc, dmin);
        __syncthreads();  // everyone is done with stage[buf] before it is refilled two tiles later
    }
}

// LIST: samples[list[i]], i < *nlist_dev, as in k_eval_sphere2
// STRIDE (LIST only): the CTAs of a fixed small grid walk the list in strides (the hand-over list, normally empty)
template <int S, int MINB, int UNROLL, bool LIST = false, bool STRIDE = false>  // S samples per thread (even), MINB resident CTAs per SM
__global__ void __launch_bounds__(EVAL_THREADS, MINB) k_eval_point3(const VrSample* __restrict__ samples, uint64_t nsamples,
                                                              const uint32_t* __restrict__ list,
                                                              const unsigned long long* __restrict__ nlist_begin_dev,
                                                              const unsigned long long* __restrict__ nlist_dev,
                                                              const float4* __restrict__ lights2,
                                                              const unsigned long long* __restrict__ nlights_dev,
                                                              float* __restrict__ out, unsigned long long* __restrict__ redo_count) {
    constexpr int P = S / 2;
    if (LIST) {  // entries [*nlist_begin_dev (0 if NULL), *nlist_dev) of the list
        const uint64_t lbeg = nlist_begin_dev ? *nlist_begin_dev : 0;
        list += lbeg;
        nsamples = *nlist_dev - lbeg;
        if ((uint64_t)blockIdx.x * (EVAL_THREADS * S) >= nsamples) return;
    }
    const uint64_t nlights = *nlights_dev;
    __shared__ __align__(128) float4 stage[2][2 * EVAL2_TILE];
    __shared__ __align__(8) uint64_t bar[2];
    const uint32_t tid = threadIdx.x;
    uint64_t base = (uint64_t)blockIdx.x * (EVAL_THREADS * S);
    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t it = 0;
    do {
    float2 nsx[P], nsy[P], nsz[P], acc[P];
    float sw[S];
#pragma unroll
    for (int p = 0; p < P; p++) {
        // samples beyond the end sit far away (d2 ~ 3e12): they never trip the d2 check and are never stored
        float4 a = make_float4(1.0e6f, 1.0e6f, 1.0e6f, 0.f), b = a;
        const uint64_t i0 = base + (uint64_t)(2 * p) * EVAL_THREADS + tid, i1 = i0 + EVAL_THREADS;
        if (i0 < nsamples) a = __ldg(reinterpret_cast<const float4*>(samples) + (LIST ? (uint64_t)list[i0] : i0));
        if (i1 < nsamples) b = __ldg(reinterpret_cast<const float4*>(samples) + (LIST ? (uint64_t)list[i1] : i1));
        nsx[p] = make_float2(-a.x, -b.x), nsy[p] = make_float2(-a.y, -b.y), nsz[p] = make_float2(-a.z, -b.z);
        sw[2 * p] = a.w, sw[2 * p + 1] = b.w;
        acc[p] = make_float2(0.f, 0.f);
    }
    float dmin = 3.0e38f;
    eval_pass_packed<P, false, UNROLL>(stage, bar, lights2, nlights, it, nsx, nsy, nsz, acc, dmin);
    if (__syncthreads_or(!(dmin >= 0.0001f))) {  // also true for a NaN d2 (non-finite input)
        if (tid == 0 && redo_count) atomicAdd(redo_count, 1ull);
#pragma unroll
        for (int p = 0; p < P; p++) acc[p] = make_float2(0.f, 0.f);
        eval_pass_packed<P, true, 2>(stage, bar, lights2, nlights, it, nsx, nsy, nsz, acc, dmin);
    }
#pragma unroll
    for (int p = 0; p < P; p++) {
        const uint64_t i0 = base + (uint64_t)(2 * p) * EVAL_THREADS + tid, i1 = i0 + EVAL_THREADS;
        if (i0 < nsamples) out[LIST ? (uint64_t)list[i0] : i0] = sw[2 * p] * acc[p].x;
        if (i1 < nsamples) out[LIST ? (uint64_t)list[i1] : i1] = sw[2 * p + 1] * acc[p].y;
    }
    base += (uint64_t)gridDim.x * (EVAL_THREADS * S);
    } while (STRIDE && base < nsamples);
}

// ---- ray-segment variant: eight samples of one camera ray share the per-light geometry ----------------------------------
// The samples of one pixel lie on a line: p_k = A + s_k * D + eps_k, A = the segment's first sample, D = the unit world
// direction of the camera ray, |eps_k| = the FP32 rounding of the sample position (~1e-5 world units on the bench scene).
// For a light at L, with M = L - A, b = M.D and h2 = |M - b D|^2 (the squared distance of the light from the ray),
//     |L - p_k|^2 = (s_k - b)^2 + h2                                   (exact for eps_k = 0)
// so once b and h2 are known (12 operations per segment and light) every further sample of the segment costs one add and
// one fma instead of the six operations of the direct form: 3.5 FMA-pipe cycles per evaluation instead of 7. With the
// distance this cheap the MUFU.RCP rate would bind alone (8 XU cycles per warp-evaluation), so two lights share one
// reciprocal, W0/a + W1/b = (W0*b + W1*a) * rcp(a*b): FMA pipe 5.5 cycles, XU pipe 4.
// Error: the model drops eps_k, a relative error of 2*|eps_k|/d on the term of a light at distance d. Segments that come
// closer than rho to any light are therefore found beforehand (k_flag_near_segments, a geometric test from the light's
// side: no per-evaluation bookkeeping in the hot loop) and handed to the exact kernel (k_eval_point3<LIST>, which also applies
// the d2 < 1e-4 guard of common_functions.h:190), as is every segment whose measured residual
// err = max_k |p_k - A - s_k D| + 3e-7 * max_k |s_k| exceeds rho / 5e3. Every term that is kept then has a relative
// error below 2*err/rho <= 4e-4, typically 1e-5 (measured in tests/test_gpu_parity.py).
// Samples that do not fill a segment (count mod 8 per pixel) go to the exact kernel as well.
__global__ void __launch_bounds__(256) k_build_segments(const uint32_t* __restrict__ counts, const unsigned long long* __restrict__ offsets,
                                                        uint64_t npix, VrSegLists sl, int pairs) {
    const uint64_t p = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    const uint32_t c = p < npix ? counts[p] : 0u;
    const uint32_t off = p < npix ? (uint32_t)offsets[p] : 0u;
    seg_build_warp(c, off, p, p < npix, sl, pairs, threadIdx.x & 31);
}

// Which segments come closer than rho to a light? Decided geometrically, from the light's side, before the evaluation:
// one thread block per entry of the light list. A camera ray t*D (D.z > 0, the reference camera looks down +z without
// rotation, point_compute_color.comp:23-33) that passes within rho of v = L - camera has a point q with |q - v| <= rho,
// so its image-plane coordinates x = q.x/q.z, y = q.y/q.z lie in the box spanned by (v.x -+ rho)/(v.z -+ rho); the pixels
// of that box are tested exactly (distance of the light from the ray, position of the foot point against the span of
// each of the pixel's segments). rho already carries a safety margin. A light whose ball reaches the camera plane (the
// camera sits inside or beside the medium) has no bounded box on that side: the near plane is clamped to a tiny positive
// depth, the box then covers the tile in that direction and the exact per-pixel test decides. Only an absurdly distant
// light makes the kernel give up and flag everything (flag_all). Flags are what k_eval_seg reads to hand a segment to
// the exact kernel.
#if !VR_CAMERA_UNROTATED_PLUS_Z
#error "k_flag_near_segments inverts vr_camera_ray's unrotated +z projection: rework its pixel box before changing the camera"
#endif
__global__ void __launch_bounds__(128) k_flag_near_segments(const VrEvalLight* __restrict__ lights, const unsigned long long* __restrict__ nlights_dev,
                                                            vr_ubo u, float tan_scale, vr_tile tile, const uint32_t* __restrict__ counts,
                                                            const unsigned long long* __restrict__ offsets,
                                                            const VrSample* __restrict__ samples, VrSegLists sl, float rho) {
    // start of an evaluation pass (block 0, before k_list_flagged appends): the exact kernel's list restarts from the
    // leftovers, nothing handed over yet (the give-up flag was zeroed with the work counters)
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) {
        sl.counts[2] = sl.counts[1];
        sl.counts[7] = 0;
    }
    const uint64_t nl = *nlights_dev;
    // gridDim.y blocks share a light's pixel box (a light inside the frustum can cover thousands of pixels of the tile)
    for (uint64_t li = blockIdx.x; li < nl; li += gridDim.x) {
    const VrEvalLight L = lights[li];
    const float cx = u.cameraPos[0], cy = u.cameraPos[1], cz = u.cameraPos[2];
    const float vx = L.x - cx, vy = L.y - cy, vz = L.z - cz;
    if (!(vx * vx + vy * vy + vz * vz <= 1.0e10f)) {  // absurdly far (or NaN): products of squared distances may leave FP32
        if (threadIdx.x == 0) *sl.flag_all = 1u;
        continue;
    }
    if (vz + rho <= 0.0f) continue;  // behind the camera plane by more than rho: no forward ray comes near
    const float W = (float)u.framebufferDim[0], H = (float)u.framebufferDim[1], aspect = W / H;
    // points of the ball that a forward ray can reach have depth in (0, vz + rho]
    const float zn = fmaxf(vz - rho, 1.0e-6f * rho), zf = vz + rho;
    const float xa = (vx - rho) / ((vx - rho) < 0.0f ? zn : zf), xb = (vx + rho) / ((vx + rho) < 0.0f ? zf : zn);
    const float ya = (vy - rho) / ((vy - rho) < 0.0f ? zn : zf), yb = (vy + rho) / ((vy + rho) < 0.0f ? zf : zn);
    // x = ((2 (px + 0.5) / W) - 1) * aspect * tan  <=>  px = (x / (aspect * tan) + 1) * W / 2 - 0.5 ; y decreases with py
    const float fx0 = (xa / (aspect * tan_scale) + 1.0f) * 0.5f * W - 0.5f, fx1 = (xb / (aspect * tan_scale) + 1.0f) * 0.5f * W - 0.5f;
    const float fy0 = (1.0f - yb / tan_scale) * 0.5f * H - 0.5f, fy1 = (1.0f - ya / tan_scale) * 0.5f * H - 0.5f;
    const float lox = (float)tile.x0, hix = (float)(tile.x0 + tile.w) - 1.0f, loy = (float)tile.y0, hiy = (float)(tile.y0 + tile.h) - 1.0f;
    if (fx1 < lox - 1.0f || fx0 > hix + 1.0f || fy1 < loy - 1.0f || fy0 > hiy + 1.0f) continue;
    const int32_t px0 = (int32_t)fmaxf(floorf(fx0) - 1.0f, lox), px1 = (int32_t)fminf(ceilf(fx1) + 1.0f, hix);
    const int32_t py0 = (int32_t)fmaxf(floorf(fy0) - 1.0f, loy), py1 = (int32_t)fminf(ceilf(fy1) + 1.0f, hiy);
    const int32_t bw = px1 - px0 + 1, bh = py1 - py0 + 1;
    if (bw <= 0 || bh <= 0) continue;
    const float rho2 = rho * rho;
    for (int64_t i = threadIdx.x + (int64_t)blockIdx.y * blockDim.x; i < (int64_t)bw * bh; i += (int64_t)blockDim.x * gridDim.y) {
        const int32_t px = px0 + (int32_t)(i % bw), py = py0 + (int32_t)(i / bw);
        const float x = ((2.0f * ((float)px + 0.5f)) / W - 1.0f) * aspect * tan_scale;
        const float y = (1.0f - (2.0f * ((float)py + 0.5f)) / H) * tan_scale;
        const float inv = rsqrtf(x * x + y * y + 1.0f);
        const float dx = x * inv, dy = y * inv, dz = inv;
        const float b = vx * dx + vy * dy + vz * dz;
        const float ex = vx - b * dx, ey = vy - b * dy, ez = vz - b * dz;
        if (ex * ex + ey * ey + ez * ez >= rho2) continue;
        const uint64_t pix = (uint64_t)(py - (int32_t)tile.y0) * tile.w + (uint32_t)(px - (int32_t)tile.x0);
        const uint32_t nseg = counts[pix] >> 3;
        if (!nseg) continue;
        const uint32_t off = (uint32_t)offsets[pix], sb = sl.seg_base[pix];
        // the segments are consecutive along the ray: stop behind the foot point (four at a time, loads issued together)
        for (uint32_t k0 = 0; k0 < nseg; k0 += 4) {
            float4 p0[4], p1[4];
#pragma unroll
            for (uint32_t q = 0; q < 4; q++) {
                const uint32_t k = k0 + q < nseg ? k0 + q : nseg - 1;
                p0[q] = __ldg(reinterpret_cast<const float4*>(samples) + off + 8u * k);
                p1[q] = __ldg(reinterpret_cast<const float4*>(samples) + off + 8u * k + 7u);
            }
            bool beyond = false;
#pragma unroll
            for (uint32_t q = 0; q < 4; q++) {
                if (k0 + q >= nseg) break;
                const float s0 = (p0[q].x - cx) * dx + (p0[q].y - cy) * dy + (p0[q].z - cz) * dz;
                const float s1 = (p1[q].x - cx) * dx + (p1[q].y - cy) * dy + (p1[q].z - cz) * dz;
                VR_CHECK(sb + k0 + q < sl.cap_flags, VR_CHK_FLAGS);
                if (b > s0 - rho && b < s1 + rho) sl.flags[sb + k0 + q] = 1;
                beyond = s0 - rho >= b;
            }
            if (beyond) break;
        }
    }
    }
}

// the flagged entries of the segment tables go onto the exact kernel's list (behind the leftovers), before the evaluation
// starts: the exact kernel can then run next to the segment kernel instead of after it. One launch for both tables: the
// first blocks16 thread blocks walk the 16-sample table, the rest the 8-sample table.
__global__ void __launch_bounds__(256) k_list_flagged(VrSegLists sl, uint32_t blocks16) {
    const bool t16 = blockIdx.x < blocks16;
    const uint64_t nseg = t16 ? sl.counts[4] : sl.counts[0];
    const uint64_t e = (uint64_t)(t16 ? blockIdx.x : blockIdx.x - blocks16) * 256 + threadIdx.x;
    if (e >= nseg) return;
    const uint32_t fi = (t16 ? sl.seg16_flag : sl.seg_flag)[e];
    VR_CHECK(fi + (t16 ? 1u : 0u) < sl.cap_flags, VR_CHK_FLAGS);
    if (!(*sl.flag_all != 0u || sl.flags[fi] != 0 || (t16 && sl.flags[fi + 1] != 0))) return;
    const uint32_t first = (t16 ? sl.seg16_first : sl.seg_first)[e];
    const uint32_t len = t16 ? 16u : 8u;
    const unsigned long long slot = atomicAdd(&sl.counts[2], (unsigned long long)len);
    for (uint32_t k = 0; k < len; k++)
        if (slot + k < sl.list_capacity) sl.list[slot + k] = first + k;
}
// start of an evaluation pass without full segments (else k_flag_near_segments does this): the exact kernel's list restarts
// from the leftovers, nothing handed over yet
__global__ void k_seg_begin(VrSegLists sl) {
    sl.counts[2] = sl.counts[1];
    sl.counts[7] = 0;
}

__device__ __forceinline__ float2 neg2(const float2 v) { return make_float2(-v.x, -v.y); }  // folds into the operand modifier

// b = (L - A).D and h2 = |(L - A) - b D|^2 for the two segments of the thread (x and y halves) against one light
__device__ __forceinline__ void seg_light_geometry(const float4 qa, const float4 qb, const float2 (&A)[3], const float2 (&D)[3],
                                                   float2& nb, float2& h2) {
    const float2 Mx = __fadd2_rn(make_float2(qa.x, qa.y), neg2(A[0]));
    const float2 My = __fadd2_rn(make_float2(qa.z, qa.w), neg2(A[1]));
    const float2 Mz = __fadd2_rn(make_float2(qb.x, qb.y), neg2(A[2]));
    float2 b = __fmul2_rn(Mx, D[0]);
    b = __ffma2_rn(My, D[1], b);
    b = __ffma2_rn(Mz, D[2], b);
    nb = neg2(b);
    const float2 px = __ffma2_rn(nb, D[0], Mx), py = __ffma2_rn(nb, D[1], My), pz = __ffma2_rn(nb, D[2], Mz);
    h2 = __fmul2_rn(px, px);
    h2 = __ffma2_rn(py, py, h2);
    h2 = __ffma2_rn(pz, pz, h2);
}

constexpr int SEG_LEN = 8;                        // samples per segment (the unit of the tables and flags)

// squared distance that enters the reciprocal. Point-type lists: the distance to the light itself. Sphere-type lists
// (sphere lights, beam sub-points; common_functions.h:182-201): the light acts from the nearest point of its sphere of
// radius R, which is |p - O| - R away from the sample; k_flag_near_segments is run with rho + R for these lists, so the
// near-surface samples, where the shader's own rounding of that point matters, are the exact kernel's.
template <bool SPHERE>
__device__ __forceinline__ float2 seg_dist2(const float2 l2, const float2 nR) {
    if (!SPHERE) return l2;
    const float2 len = __fmul2_rn(l2, make_float2(rsqrt_approx(l2.x), rsqrt_approx(l2.y)));
    const float2 d = __fadd2_rn(len, nR);
    return __fmul2_rn(d, d);
}

// LEN = 8: one segment per half; LEN = 16: two consecutive segments of the same pixel per half (table sl.seg16_*), so the
// per-light geometry is shared by 16 samples; their weights are read again at the end instead of held in registers.
//
// Work distribution. A chunk is 2 * THREADS table entries, and every sample's light sum is DEFINED as the sum over
// SEG_SPLIT consecutive slices of the light list, added in slice order and weighted once: out = w*(((S0 + S1) + S2) + S3).
// A "full" unit is one CTA that walks the whole list for a chunk in one uninterrupted loop (entries stay in registers); at
// a slice boundary it moves its accumulators into a running total in shared memory (3 x 16 LDS/STS per thread per chunk,
// no global traffic, no pipeline restart). A "slice" unit is one CTA per (chunk, slice): it writes its slice's sums to a
// scratch buffer, and whichever of a chunk's SEG_SPLIT units finishes last (a per-chunk counter) adds the four partial
// sums in slice order and stores the results -- nobody waits for anybody. The full units (instance SLICE = false, main
// stream) cover all chunks but the last `slots` (one resident wave); those last chunks run as slice units (instance
// SLICE = true, a lower-priority stream: its CTAs take the SM slots the full units free). So the end of the pass is made
// of units a quarter as long, and the GPU drains evenly: the tail of a launch whose chunks fill 3.05 waves (one of eight
// GPUs' bands) costs a quarter of a CTA time instead of a whole one. The result does not depend on which kind of unit
// produced it -- a tiled render equals the whole frame bit for bit (tests/test_gpu_configs.py).
constexpr int SEG_SPLIT = 4;
// lights per slice: a quarter of the list rounded up to whole pairs (the loop shares one reciprocal between two lights, so
// slice boundaries must not split the pairs the unsliced list would form), at least 64
__host__ __device__ __forceinline__ uint64_t seg_slice_len(uint64_t nlights) {
    const uint64_t q = ((nlights + SEG_SPLIT - 1) / SEG_SPLIT + 1) & ~1ull;
    return q < 64 ? 64 : q;
}
// dynamic shared memory of k_eval_seg: the two light stages, the two mbarriers, the running totals of a full unit
constexpr uint32_t SEG_STAGE_BYTES = 2u * 2u * EVAL2_TILE * 16u;
constexpr uint32_t seg_smem_bytes(int len, int threads) { return SEG_STAGE_BYTES + 128u + (uint32_t)len * (uint32_t)threads * 8u; }

// SLICE = false: the full units (chunk = blockIdx.x, all bookkeeping uniform: the loop keeps every register it had before
// the slicing existed); SLICE = true: the slice units of the last `slots` chunks, taken from the ticket counter.
template <int THREADS, int MINB, bool SPHERE, int LEN, bool SLICE>
__global__ void __launch_bounds__(THREADS, MINB) k_eval_seg(VrGridDev g, vr_ubo u, float tan_scale, vr_tile tile,
                                                                  const VrSample* __restrict__ samples, VrSegLists sl,
                                                                  const float4* __restrict__ lights2,
                                                                  const unsigned long long* __restrict__ nlights_dev, float rho,
                                                                  float radius, float* __restrict__ out, uint32_t slots) {
    extern __shared__ __align__(128) unsigned char seg_smem[];
    float4 (*stage)[2 * EVAL2_TILE] = reinterpret_cast<float4 (*)[2 * EVAL2_TILE]>(seg_smem);
    uint64_t* bar = reinterpret_cast<uint64_t*>(seg_smem + SEG_STAGE_BYTES);
    float2* part = reinterpret_cast<float2*>(seg_smem + SEG_STAGE_BYTES + 128u);  // [k][tid]: conflict-free 8-byte accesses (full units)
    constexpr int SEG_PER_CTA = 2 * THREADS;  // two entries per thread: the x and y halves of every FP32x2 register
    const float2 nR = make_float2(-radius, -radius);
    constexpr bool KEEP_W = LEN <= 8;
    const uint64_t nseg = LEN == 16 ? sl.counts[4] : sl.counts[0];
    const uint32_t* __restrict__ tab_first = LEN == 16 ? sl.seg16_first : sl.seg_first;
    const uint32_t* __restrict__ tab_pix = LEN == 16 ? sl.seg16_pix : sl.seg_pix;
    const uint32_t* __restrict__ tab_flag = LEN == 16 ? sl.seg16_flag : sl.seg_flag;
    uint32_t* __restrict__ sched = LEN == 16 ? sl.sched16 : sl.sched8;  // [0] ticket counter, [1 + c] slices of chunk c done
    const uint32_t tid = threadIdx.x;
    const uint32_t nchunks = (uint32_t)((nseg + SEG_PER_CTA - 1) / SEG_PER_CTA);
    const uint32_t nsliced = nchunks < slots ? nchunks : slots;
    const uint32_t nfull = nchunks - nsliced;
    uint32_t chunk = blockIdx.x, q0 = 0;
    constexpr bool sliced = SLICE;
    if (SLICE) {
        chunk = nfull + blockIdx.x / SEG_SPLIT, q0 = blockIdx.x % SEG_SPLIT;
        if (chunk >= nchunks) return;  // the launch is sized by the capacity, the segment count lives on the device
    } else if (chunk >= nfull) {
        return;
    }
    uint32_t* __restrict__ done = sched + 1 + chunk;
    VR_CHECK(1 + chunk < sl.sched_words, VR_CHK_SCHED);
    const uint64_t seg_base = (uint64_t)chunk * SEG_PER_CTA;
    const bool redo_all = *sl.flag_all != 0u;
    bool idle;
    {
        // a CTA whose entries are all on the exact kernel's list already (near a light, or flag_all) has nothing to evaluate
        bool listed = true;
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const uint64_t seg = seg_base + tid + (uint64_t)h * THREADS;
            if (seg < nseg && !redo_all) {
                const uint32_t fi = tab_flag[seg];
                listed = listed && (sl.flags[fi] != 0 || (LEN == 16 && sl.flags[fi + 1] != 0));
            }
        }
        idle = __syncthreads_and(listed) != 0;
    }
    if (idle) return;  // (all SEG_SPLIT slice units of a chunk see the same flags: none of them reports to the chunk's counter)
    // (list lengths are below 2^32: vr_dispatch_compute_color refuses longer expansions; 32-bit counters spare registers)
    const uint32_t nlights = (uint32_t)*nlights_dev;
    const uint32_t slice_len = (uint32_t)seg_slice_len(nlights);
    // the lights this unit walks: the whole list, or slice q0 of it
    const uint32_t lo = sliced ? (q0 * slice_len < nlights ? q0 * slice_len : nlights) : 0u;
    const uint32_t nl = !sliced ? nlights : (nlights - lo < slice_len ? nlights - lo : slice_len);
    const float4* __restrict__ src = lights2 + 2 * (uint64_t)lo;
    VR_CHECK(lo + nl <= nlights, VR_CHK_LIGHTS);
    const uint32_t ntiles = (nl + EVAL2_TILE - 1) / EVAL2_TILE;
    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if (ntiles > 0) {  // the first tile travels while the entries are loaded
            const uint32_t cnt = nl < EVAL2_TILE ? nl : EVAL2_TILE;
            mbar_expect_tx(&bar[0], cnt * 32u);
            tma_bulk_g2s(&stage[0][0], src, cnt * 32u, &bar[0]);
        }
    }
    __syncthreads();

    float2 s[LEN], acc[LEN], A[3], D[3];
    float sw[2][KEEP_W ? LEN : 1];
    uint32_t first[2];
    bool redo[2] = {false, false};  // per segment, so that what the exact kernel evaluates does not depend on the pairing
    bool skip[2] = {false, false};  // on the exact kernel's list already: that kernel (another stream) owns the result
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const uint64_t seg = seg_base + tid + (uint64_t)h * THREADS;
        float a[3] = {1.0e6f, 1.0e6f, 1.0e6f}, d[3] = {1.0f, 0.0f, 0.0f}, sv[LEN];
        first[h] = 0xffffffffu;
#pragma unroll
        for (int k = 0; k < LEN; k++) {
            sv[k] = 0.0f;
            if (KEEP_W) sw[h][k] = 0.0f;
        }
        if (seg < nseg) {
            first[h] = tab_first[seg];
            VR_CHECK((uint64_t)first[h] + LEN <= sl.nslots, VR_CHK_SLOTS);
            const uint32_t pix = tab_pix[seg];
            const VrCamRay r = vr_camera_ray(g, u, tan_scale, tile.x0 + pix % tile.w, tile.y0 + pix / tile.w);
            // world direction of the ray: the linear part of the index->world map applied to the index-space direction
            const float vx = r.d.x * g.mat[0] + r.d.y * g.mat[1] + r.d.z * g.mat[2];
            const float vy = r.d.x * g.mat[3] + r.d.y * g.mat[4] + r.d.z * g.mat[5];
            const float vz = r.d.x * g.mat[6] + r.d.y * g.mat[7] + r.d.z * g.mat[8];
            const float inv = 1.0f / sqrtf(vx * vx + vy * vy + vz * vz);
            d[0] = vx * inv, d[1] = vy * inv, d[2] = vz * inv;
            const float4* sp = reinterpret_cast<const float4*>(samples) + first[h];
            float res2 = 0.0f, smax = 0.0f;
#pragma unroll
            for (int k = 0; k < LEN; k++) {
                const float4 q = __ldg(sp + k);
                if (k == 0) a[0] = q.x, a[1] = q.y, a[2] = q.z;
                const float ex = q.x - a[0], ey = q.y - a[1], ez = q.z - a[2];
                const float sk = ex * d[0] + ey * d[1] + ez * d[2];
                const float rx = ex - sk * d[0], ry = ey - sk * d[1], rz = ez - sk * d[2];
                res2 = fmaxf(res2, rx * rx + ry * ry + rz * rz);
                smax = fmaxf(smax, fabsf(sk));
                sv[k] = sk;
                if (KEEP_W) sw[h][k] = q.w;
            }
            // Segments near a light were put on the exact kernel's list by k_list_flagged already. What is decided here: samples
            // too far off their line for the bound to hold -> hand the entry over as well (rare: a safety net for grid scales
            // where FP32 positions are coarse).
            const uint32_t fi = tab_flag[seg];  // the pixel's segments are numbered consecutively for the flags
            const bool listed = redo_all || sl.flags[fi] != 0 || (LEN == 16 && sl.flags[fi + 1] != 0);
            skip[h] = listed;
            redo[h] = !listed && !(5.0e3f * (sqrtf(res2) + 3.0e-7f * smax) <= rho);
        }
#pragma unroll
        for (int k = 0; k < LEN; k++) {
            if (h == 0) s[k].x = sv[k];
            else s[k].y = sv[k];
        }
#pragma unroll
        for (int c = 0; c < 3; c++) {
            if (h == 0) A[c].x = a[c], D[c].x = d[c];
            else A[c].y = a[c], D[c].y = d[c];
        }
    }

#pragma unroll
    for (int k = 0; k < LEN; k++) acc[k] = make_float2(0.f, 0.f);
    // a full unit moves its accumulators into the running total at every slice boundary
    uint32_t boundary = sliced ? 0xffffffffu : slice_len;
    bool spilled = false;
    for (uint32_t tl = 0; tl < ntiles; tl++) {
        const uint32_t buf = tl & 1u;
        if (tid == 0 && tl + 1 < ntiles) {
            const uint32_t nxt = (tl + 1) * EVAL2_TILE;
            const uint32_t cnt = (nl - nxt) < EVAL2_TILE ? (nl - nxt) : EVAL2_TILE;
            mbar_expect_tx(&bar[buf ^ 1], cnt * 32u);
            tma_bulk_g2s(&stage[buf ^ 1][0], src + 2 * (uint64_t)nxt, cnt * 32u, &bar[buf ^ 1]);
        }
        mbar_wait(&bar[buf], (tl >> 1) & 1u);
        const uint32_t tlo = tl * EVAL2_TILE;
        const int cnt = (int)((nl - tlo) < EVAL2_TILE ? (nl - tlo) : EVAL2_TILE);
        const float4* __restrict__ L = stage[buf];
        int l = 0;
        while (l < cnt) {
            const int end = boundary - tlo < (uint32_t)cnt ? (int)(boundary - tlo) : cnt;  // (boundary > tlo + l always)
#pragma unroll 1
            for (; l + 2 <= end; l += 2) {
                const float4 qa = L[2 * l], qb = L[2 * l + 1], ra = L[2 * l + 2], rb = L[2 * l + 3];
                float2 nb0, h0, nb1, h1;
                seg_light_geometry(qa, qb, A, D, nb0, h0);
                seg_light_geometry(ra, rb, A, D, nb1, h1);
                const float2 W0 = make_float2(qb.z, qb.w), W1 = make_float2(rb.z, rb.w);
#pragma unroll
                    for (int k = 0; k < LEN; k++) {
                        const float2 u0 = __fadd2_rn(s[k], nb0), u1 = __fadd2_rn(s[k], nb1);
                        const float2 a = seg_dist2<SPHERE>(__ffma2_rn(u0, u0, h0), nR), b = seg_dist2<SPHERE>(__ffma2_rn(u1, u1, h1), nR);
                        const float2 ab = __fmul2_rn(a, b);
                        const float2 num = __ffma2_rn(W1, a, __fmul2_rn(W0, b));
                        acc[k] = __ffma2_rn(num, make_float2(rcp_approx(ab.x), rcp_approx(ab.y)), acc[k]);
                    }
            }
            if (l < end) {  // odd tail of the list (slices are whole pairs: this is the list's last light)
                const float4 qa = L[2 * l], qb = L[2 * l + 1];
                float2 nb0, h0;
                seg_light_geometry(qa, qb, A, D, nb0, h0);
                const float2 W0 = make_float2(qb.z, qb.w);
#pragma unroll
                for (int k = 0; k < LEN; k++) {
                    const float2 u0 = __fadd2_rn(s[k], nb0);
                    const float2 a = seg_dist2<SPHERE>(__ffma2_rn(u0, u0, h0), nR);
                    acc[k] = __ffma2_rn(W0, make_float2(rcp_approx(a.x), rcp_approx(a.y)), acc[k]);
                }
                l++;
            }
            if (tlo + (uint32_t)l == boundary && boundary < nl) {  // end of a slice with more to come: total (+)= acc
#pragma unroll
                for (int k = 0; k < LEN; k++) {
                    float2* pp = part + k * THREADS + tid;
                    *pp = spilled ? __fadd2_rn(*pp, acc[k]) : acc[k];
                    acc[k] = make_float2(0.f, 0.f);
                }
                spilled = true;
                boundary += slice_len;
            }
        }
        __syncthreads();  // everyone is done with stage[buf] before it is refilled two tiles later
    }
    if (spilled) {
#pragma unroll
        for (int k = 0; k < LEN; k++) acc[k] = __fadd2_rn(part[k * THREADS + tid], acc[k]);
    }

    // ---- full unit: out = w * total
    if (!SLICE) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
            if (first[h] == 0xffffffffu || skip[h]) continue;
#pragma unroll
            for (int k = 0; k < LEN; k++) {
                const float v = h == 0 ? acc[k].x : acc[k].y;
                out[first[h] + k] = (KEEP_W ? sw[h][k] : __ldg(&samples[first[h] + k].w)) * v;
            }
        }
    } else {
        // ---- slice unit: partial sums to the scratch buffer ([slice][k][entry of the chunk]: coalesced); the unit that
        // finishes last adds the chunk's SEG_SPLIT partial sums in slice order
        float* __restrict__ pc = (LEN == 16 ? sl.part16 : sl.part8) + (uint64_t)(chunk - nfull) * (SEG_SPLIT * LEN * SEG_PER_CTA);
        float* __restrict__ pq = pc + q0 * (LEN * SEG_PER_CTA);
        VR_CHECK((uint64_t)(chunk - nfull + 1) * (SEG_SPLIT * LEN * SEG_PER_CTA) <= (LEN == 16 ? sl.part16_floats : sl.part8_floats), VR_CHK_PART);
#pragma unroll
        for (int h = 0; h < 2; h++) {
            if (first[h] == 0xffffffffu || skip[h]) continue;
#pragma unroll
            for (int k = 0; k < LEN; k++) __stcg(pq + k * SEG_PER_CTA + h * THREADS + tid, h == 0 ? acc[k].x : acc[k].y);
        }
        __threadfence();
        __shared__ uint32_t s_arrived;
        __syncthreads();
        if (tid == 0) s_arrived = atomicAdd(done, 1u);
        __syncthreads();
        if (s_arrived == SEG_SPLIT - 1) {
            __threadfence();
#pragma unroll
            for (int h = 0; h < 2; h++) {
                if (first[h] == 0xffffffffu || skip[h]) continue;
#pragma unroll
                for (int k = 0; k < LEN; k++) {
                    const float* pe = pc + k * SEG_PER_CTA + h * THREADS + tid;
                    float t = __ldcg(pe);
#pragma unroll
                    for (int q = 1; q < SEG_SPLIT; q++) t = __fadd_rn(t, __ldcg(pe + q * (LEN * SEG_PER_CTA)));
                    out[first[h] + k] = (KEEP_W ? sw[h][k] : __ldg(&samples[first[h] + k].w)) * t;
                }
            }
        }
    }

    if (q0 == 0) {  // entries handed over to the exact kernel (second list): announced once per chunk
        const uint32_t npush = (redo[0] ? LEN : 0) + (redo[1] ? LEN : 0);
        unsigned long long slot = 0;
        if (npush) slot = atomicAdd(&sl.counts[7], (unsigned long long)npush);
#pragma unroll
        for (int h = 0; h < 2; h++) {
            if (!redo[h]) continue;
#pragma unroll
            for (int k = 0; k < LEN; k++)
                if (slot + k < sl.list_capacity) sl.list2[slot + k] = first[h] + k;
            slot += LEN;
        }
    }
}

// ---- opt-in physics: Henyey-Greenstein phase factor and light->sample transmittance per evaluation.
// Not a hot path: one thread per sample walks the evaluation list from global memory (every thread of a warp reads the
// same entry) and applies oracle/vr_oracle.cpp's physFactor to each term with individually rounded operations; the
// transmittance term marches the grid from every light to every sample (orders of magnitude more lookups than the
// reference's gather). Point-type lists: getPointLightContribution; sphere-type: from the nearest point of the sphere.
template <bool SPHERE>
__global__ void __launch_bounds__(128) k_eval_phys(VrGridDev g, vr_ubo u, VrPhysics phys, const VrSample* __restrict__ samples,
                                                    uint64_t nsamples, const VrEvalLight* __restrict__ lights,
                                                    const unsigned long long* __restrict__ nlights_dev, float* __restrict__ out) {
    const uint64_t idx = (uint64_t)blockIdx.x * 128 + threadIdx.x;
    if (idx >= nsamples) return;
    const uint64_t nlights = *nlights_dev;
    const float4 sv = __ldg(reinterpret_cast<const float4*>(samples) + idx);
    const vr_f3 p = vr_make3(sv.x, sv.y, sv.z);
    VrAcc acc;
    vr_acc_init(acc);
    float total = 0.0f;
    for (uint64_t l = 0; l < nlights; l++) {
        const float4 q = __ldg(reinterpret_cast<const float4*>(lights) + l);
        vr_f3 lp = vr_make3(q.x, q.y, q.z);
        if (SPHERE) lp = vr_add3(lp, vr_scale3(vr_normalize3(vr_sub3(p, lp)), u.beamRadius));  // getClosestPointOnSphere
        const vr_f3 L = vr_sub3(lp, p);
        const float d2 = vr_dot3(L, L);
        if (d2 < 0.0001f) continue;
        total = vr_add(total, vr_mul(vr_div(q.w, d2), vr_phys_factor(g, u, phys, acc, p, lp)));
    }
    out[idx] = vr_mul(sv.w, total);
}

__global__ void __launch_bounds__(128) k_trilinear(VrGridDev g, const float* __restrict__ xyz, uint64_t n, float* __restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * 128 + threadIdx.x;
    if (i >= n) return;
    VrAcc acc;
    vr_acc_init(acc);
    out[i] = vr_trilinear(g, acc, vr_make3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]));
}

// =====================================================================================================
// resolve
// =====================================================================================================
// round-half-even UNORM8 store of a [0,1] value (Vulkan rgba8 conversion)
__device__ __forceinline__ unsigned char to_unorm8(float v) {
    v = vr_clamp(v, 0.0f, 1.0f);
    if (v != v) v = 0.0f;
    return (unsigned char)__float2int_rn(vr_mul(v, 255.0f));
}
// progressive average of point_compute_color.comp:99-105 on the float accumulator and on the rgba8 image.
// nn == 0 selects the SUM mode of vr_set_accumulate_mode (sample-parallel progressive rendering): the float accumulator
// adds the frame colour, the rgba8 image is left alone (its quantised running average is sequential by construction).
__device__ __forceinline__ void accumulate_pixel(float fc, float nm1, float nn, float4* accum, uchar4* rgba8, uint64_t gp) {
    const float4 prev = accum[gp];
    float4 nw;
    if (nn == 0.0f) {
        nw.x = vr_add(prev.x, fc), nw.y = vr_add(prev.y, fc), nw.z = vr_add(prev.z, fc), nw.w = 1.0f;
        accum[gp] = nw;
        return;
    }
    nw.x = vr_div(vr_add(vr_mul(prev.x, nm1), fc), nn);
    nw.y = vr_div(vr_add(vr_mul(prev.y, nm1), fc), nn);
    nw.z = vr_div(vr_add(vr_mul(prev.z, nm1), fc), nn);
    nw.w = 1.0f;
    accum[gp] = nw;
    const uchar4 p8 = rgba8[gp];
    uchar4 n8;
    n8.x = to_unorm8(vr_div(vr_add(vr_mul(vr_div(vr_u2f(p8.x), 255.0f), nm1), fc), nn));
    n8.y = to_unorm8(vr_div(vr_add(vr_mul(vr_div(vr_u2f(p8.y), 255.0f), nm1), fc), nn));
    n8.z = to_unorm8(vr_div(vr_add(vr_mul(vr_div(vr_u2f(p8.z), 255.0f), nm1), fc), nn));
    n8.w = 255;
    rgba8[gp] = n8;
}
// frame colour of pixel gp to this GPU's buffer and to every peer's (plain stores to peer-mapped memory travel over
// NVLink; the caller orders them with a cross-GPU barrier before the peers read)
__device__ __forceinline__ void store_frame_color(const VrFrameColorOut& o, uint64_t gp, float fc) {
    o.local[gp] = fc;
    for (uint32_t q = 0; q < o.npeers; q++) o.peer[q][gp] = fc;
}
__global__ void __launch_bounds__(256) k_resolve(vr_ubo u, vr_tile tile, const uint32_t* __restrict__ counts,
                                                  const unsigned long long* __restrict__ offsets, const float* __restrict__ contrib,
                                                  const VrLightCounter* __restrict__ counter, VrFrameColorOut frame_color,
                                                  float4* __restrict__ accum, uchar4* __restrict__ rgba8, int sum_mode) {
    const uint64_t pix = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= (uint64_t)tile.w * tile.h) return;
    const uint32_t lx = (uint32_t)(pix % tile.w), ly = (uint32_t)(pix / tile.w);
    const uint64_t gp = (uint64_t)(tile.y0 + ly) * u.framebufferDim[0] + (tile.x0 + lx);
    const uint32_t n = counts[pix];
    const float* c = contrib + offsets[pix];
    float finalColor = 0.0f;
    for (uint32_t s = 0; s < n; s++) finalColor = vr_add(finalColor, c[s]);  // finalColor += dC, in march order
    float fc = vr_div(finalColor, vr_u2f(counter->lightCount));
    fc = vr_clamp(fc, 0.0f, 1.0f);
    if (fc != fc) fc = 0.0f;  // lightCount == 0 -> 0/0; a NaN stores as 0 (Appendix B)
    store_frame_color(frame_color, gp, fc);
    const float total = sum_mode ? 0.0f : vr_u2f(u.frameCount);  // float totalInvocations = frameCount * 1
    accumulate_pixel(fc, vr_sub(total, 1.0f), total, accum, rgba8, gp);
}

// running average of an already known frame colour (rows rendered by another GPU and exchanged over NVLink)
__global__ void __launch_bounds__(256) k_accumulate(vr_ubo u, vr_tile tile, const float* __restrict__ frame_color,
                                                     float4* __restrict__ accum, uchar4* __restrict__ rgba8, int sum_mode) {
    const uint64_t pix = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= (uint64_t)tile.w * tile.h) return;
    const uint32_t lx = (uint32_t)(pix % tile.w), ly = (uint32_t)(pix / tile.w);
    const uint64_t gp = (uint64_t)(tile.y0 + ly) * u.framebufferDim[0] + (tile.x0 + lx);
    const float total = sum_mode ? 0.0f : vr_u2f(u.frameCount);
    accumulate_pixel(frame_color[gp], vr_sub(total, 1.0f), total, accum, rgba8, gp);
}
// end of a sample-parallel progressive run: the float accumulator holds the SUM of `total` frame colours (this GPU's, plus
// the other GPUs' after the reduce); make it their average and derive the rgba8 image from that average
__global__ void __launch_bounds__(256) k_finish_sum(float4* __restrict__ accum, uchar4* __restrict__ rgba8, uint64_t npix, float total) {
    const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix) return;
    const float4 s = accum[p];
    const float4 a = make_float4(vr_div(s.x, total), vr_div(s.y, total), vr_div(s.z, total), 1.0f);
    accum[p] = a;
    rgba8[p] = make_uchar4(to_unorm8(a.x), to_unorm8(a.y), to_unorm8(a.z), 255);
}
// dst += src over n float4 (the reduce of the sample-parallel mode inside one process: src is a peer GPU's accumulator)
__global__ void __launch_bounds__(256) k_add_accum(float4* __restrict__ dst, const float4* __restrict__ src, uint64_t npix) {
    const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix) return;
    const float4 a = dst[p], b = src[p];
    dst[p] = make_float4(vr_add(a.x, b.x), vr_add(a.y, b.y), vr_add(a.z, b.z), 1.0f);
}
// Cross-GPU barrier inside one process (vr_group): GPU `self` announces epoch `epoch` in every GPU's flag array (peer
// stores over NVLink) and waits until all n GPUs have announced it in its own. The stores of the kernels that ran before
// this one on the stream (frame colours written to the peers' buffers by k_resolve) are ordered in front of the flag by
// the system-scope fence, so a GPU that sees the flag also sees them. A peer that never arrives (its host thread failed,
// or its stream is queued behind this very kernel) must not hang the GPU: after timeout_ns the wait gives up and leaves
// a mark in word VR_MAX_PEERS of this GPU's array, which vr_group_sync reports.
__global__ void k_peer_barrier(VrPeerFlags f, uint32_t self, uint32_t n, uint32_t epoch, unsigned long long timeout_ns) {
    const uint32_t q = threadIdx.x;
    if (q < n) {
        __threadfence_system();
        volatile uint32_t* dst = f.flags[q] + self;
        *dst = epoch;
        volatile uint32_t* mine = f.flags[self] + q;
        unsigned long long t0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        while ((int32_t)(*mine - epoch) < 0) {
            __nanosleep(64);
            unsigned long long t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > timeout_ns) {
                f.flags[self][VR_MAX_PEERS] = epoch;
                break;
            }
        }
        __threadfence_system();
    }
}

// =====================================================================================================
// path mode (shaders/path_compute_color.comp), one pixel per thread, exact op order
// =====================================================================================================
// The shader's shadow term multiplies `light` by the same factor a = exp(-v1*sigma*step) once per step between the sample
// and the light source (path_compute_color.comp:17-28; v1 is re-read from ONE voxel next to the light, :20). With
// rayMarchingStepSize <= 1 that voxel is one of the 27 around the light source, so the whole term is a function of
// (neighbour, step count): k_path_table runs the shader's multiply chain once per neighbour and k_path looks the value up.
// Same recurrence, same rounding -> bit-identical to looping per sample; the loop remains as the fallback.
__global__ void k_path_table(VrGridDev g, vr_ubo u, uint32_t nmax, float* __restrict__ tab) {
    const uint32_t nb = threadIdx.x;
    if (nb >= 27) return;
    const vr_f3 ls = vr_world_to_index(g, vr_make3(u.lightSourceWorldPos[0], u.lightSourceWorldPos[1], u.lightSourceWorldPos[2]));
    const int32_t i = (int32_t)vr_floor(ls.x) + (int32_t)(nb / 9) - 1, j = (int32_t)vr_floor(ls.y) + (int32_t)((nb / 3) % 3) - 1,
                  k = (int32_t)vr_floor(ls.z) + (int32_t)(nb % 3) - 1;
    VrAcc acc;
    vr_acc_init(acc);
    const float v1 = vr_lookup(g, acc, i, j, k);
    const float a = vr_exp(vr_mul(vr_mul(-v1, u.absorptionCoefficient), u.rayMarchingStepSize));
    float light = vr_div(u.photonInitialIntensity, 10000.0f);
    float* row = tab + (size_t)nb * (nmax + 1);
    row[0] = light;
    for (uint32_t s = 1; s <= nmax; s++) {
        light = vr_mul(light, a);
        row[s] = light;
    }
}

__global__ void __launch_bounds__(128) k_path(VrGridDev g, vr_ubo u, float tan_scale, vr_tile tile, VrFrameColorOut frame_color,
                                               float4* __restrict__ accum, uchar4* __restrict__ rgba8, VrWorkCounters* wc,
                                               const float* __restrict__ tab, uint32_t nmax, int sum_mode) {
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t lx = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const uint32_t ly = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
    const bool active = lx < tile.w && ly < tile.h;
    uint32_t sref = 0, seff = 0;
    if (active) {
        const uint32_t px = tile.x0 + lx, py = tile.y0 + ly;
        VrSeed seed = {px * u.frameCount, py * u.frameCount, 0u};
        const VrCamRay r = vr_camera_ray(g, u, tan_scale, px, py);
        const float step = u.rayMarchingStepSize;
        const float T = 1.0f;  // never updated by the shader
        float t = r.tmin;
        vr_f3 o = r.o, d = r.d;
        vr_f3 cur = vr_add3(o, vr_scale3(d, r.tmin));
        const vr_f3 ls = vr_world_to_index(g, vr_make3(u.lightSourceWorldPos[0], u.lightSourceWorldPos[1], u.lightSourceWorldPos[2]));
        const int32_t lsi = (int32_t)vr_floor(ls.x), lsj = (int32_t)vr_floor(ls.y), lsk = (int32_t)vr_floor(ls.z);
        const float start = vr_div(u.photonInitialIntensity, 10000.0f);
        float finalColor = 0.0f;
        VrAcc acc, acc2;
        vr_acc_init(acc);
        vr_acc_init(acc2);
        while (r.hit && t < r.tmax && T > 0.001f && u.photonInitialIntensity > 0.01f) {
            const int32_t i = (int32_t)vr_floor(cur.x), j = (int32_t)vr_floor(cur.y), k = (int32_t)vr_floor(cur.z);
            if (vr_in_bbox(g, i, j, k)) {
                const float val = vr_lookup(g, acc, i, j, k);
                sref++;
                if (val > 0.0f) {
                    seff++;
                    const float roll = vr_randf_inc(seed);
                    if (roll < u.scatteringProbability) {
                        const float q1 = vr_randf_inc(seed);
                        const float q2 = vr_randf_inc(seed);
                        d = vr_random_dir(q1, q2);
                        o = cur;
                        t = 0.0f;
                    }
                    // computeLightContribution(currentPosLocal, photonInitialIntensity), path_compute_color.comp:9-31
                    const vr_f3 rl = vr_sub3(ls, cur);
                    const vr_f3 rd = vr_normalize3(rl);
                    const float q = vr_div(vr_length3(rl), step);
                    const uint32_t steps = (q >= 1.0f) ? (q < 4.0e9f ? (uint32_t)q : 0xFFFFFFFFu) : 0u;
                    float light = start;
                    if (steps > 0) {
                        // the loop body reads the same voxel every iteration (:20): hoist the lookup, keep the multiplies
                        const vr_f3 c2 = vr_add3(ls, vr_scale3(rd, step));
                        const int32_t ci = (int32_t)vr_floor(c2.x), cj = (int32_t)vr_floor(c2.y), ck = (int32_t)vr_floor(c2.z);
                        const uint32_t di = (uint32_t)(ci - lsi + 1), dj = (uint32_t)(cj - lsj + 1), dk = (uint32_t)(ck - lsk + 1);
                        if (tab && di < 3u && dj < 3u && dk < 3u && steps <= nmax) {
                            light = __ldg(tab + (size_t)(di * 9u + dj * 3u + dk) * (nmax + 1) + steps);
                        } else {
                        const float v1 = vr_lookup(g, acc2, ci, cj, ck);
                        const float a = vr_exp(vr_mul(vr_mul(-v1, u.absorptionCoefficient), step));
                        if (a == 1.0f) {
                            // x * 1 == x: nothing to do
                        } else {
                            for (uint32_t s = 0; s < steps; s++) {
                                light = vr_mul(light, a);
                                if (light == 0.0f) break;  // 0 * a stays 0 (a is finite and non-negative)
                            }
                        }
                        }
                    }
                    finalColor = vr_add(finalColor, vr_mul(vr_mul(vr_mul(T, val), step), light));
                }
            }
            t = vr_add(t, step);
            cur = vr_add3(o, vr_scale3(d, t));
        }
        float fc = vr_clamp(finalColor, 0.0f, 1.0f);
        if (fc != fc) fc = 0.0f;
        const uint64_t gp = (uint64_t)py * u.framebufferDim[0] + px;
        store_frame_color(frame_color, gp, fc);
        // (previousColor * (contributionCount - 1) + frameColor) / contributionCount with uint contributionCount
        accumulate_pixel(fc, vr_u2f(u.frameCount - 1u), sum_mode ? 0.0f : vr_u2f(u.frameCount), accum, rgba8, gp);
    }
    if (wc) {
        unsigned long long a = sref, b = seff;
#pragma unroll
        for (int o2 = 16; o2 > 0; o2 >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, o2);
            b += __shfl_xor_sync(0xffffffffu, b, o2);
        }
        if (lane == 0) {
            atomicAdd(&wc->S_ref, a);
            atomicAdd(&wc->S_eff, b);
        }
    }
}

// =====================================================================================================
// CPU_test/main.cpp:25-43, 89-126
// =====================================================================================================
__global__ void __launch_bounds__(128) k_cpu_test(VrGridDev g, uint32_t w, uint32_t h, float aspect, float scale, float* __restrict__ density) {
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t i = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const uint32_t j = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
    if (i >= w || j >= h) return;
    // float x = (2 * (i + 0.5) / width - 1) * aspectRatio * scale;   (double arithmetic, narrowed)
    const double xd = __dmul_rn(__dmul_rn(__dsub_rn(__ddiv_rn(__dmul_rn(2.0, __dadd_rn((double)(int)i, 0.5)), (double)(int)w), 1.0), (double)aspect), (double)scale);
    const double yd = __dmul_rn(__dsub_rn(1.0, __ddiv_rn(__dmul_rn(2.0, __dadd_rn((double)(int)j, 0.5)), (double)(int)h)), (double)scale);
    const float x = (float)xd, y = (float)yd;
    // glm::normalize(v) = v * inversesqrt(dot(v, v))
    const float inv = vr_div(1.0f, vr_sqrt(vr_add(vr_add(vr_mul(x, x), vr_mul(y, y)), 1.0f)));
    const float dx = vr_mul(x, inv), dy = vr_mul(y, inv), dz = vr_mul(1.0f, inv);
    VrAcc acc;
    vr_acc_init(acc);
    float t = 0.0f, dens = 0.0f;
    while (t < 1200.0f) {
        const float px = vr_add(0.0f, vr_mul(t, dx)), py = vr_add(250.0f, vr_mul(t, dy)), pz = vr_add(-800.0f, vr_mul(t, dz));
        const float v = vr_lookup(g, acc, (int32_t)vr_floor(px), (int32_t)vr_floor(py), (int32_t)vr_floor(pz));
        dens = vr_add(dens, vr_mul(v, 1.0f));
        t = vr_add(t, 1.0f);
    }
    density[(uint64_t)j * w + i] = dens;
}

// nearest-voxel lookups for a list of coordinates; each thread walks `per_thread` consecutive entries with one accessor
// so the cached-leaf / cached-lower / cached-upper / root paths are all exercised (test hook for the traversal)
__global__ void k_lookup(VrGridDev g, const int32_t* __restrict__ ijk, uint64_t n, uint32_t per_thread, float* __restrict__ out) {
    const uint64_t first = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * per_thread;
    VrAcc acc;
    vr_acc_init(acc);
    for (uint32_t q = 0; q < per_thread && first + q < n; q++) {
        const uint64_t e = first + q;
        out[e] = vr_lookup(g, acc, ijk[3 * e], ijk[3 * e + 1], ijk[3 * e + 2]);
    }
}

__global__ void k_clear(float4* accum, uchar4* rgba8, float* frame_color, uint64_t npix) {
    const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix) return;
    accum[p] = make_float4(0.f, 0.f, 0.f, 1.f);  // vkCmdClearColorImage (0,0,0,1), src/main.cpp:760
    rgba8[p] = make_uchar4(0, 0, 0, 255);
    frame_color[p] = 0.0f;
}
__global__ void __launch_bounds__(256) k_row_sums(const uint32_t* __restrict__ counts, uint32_t w, unsigned long long* __restrict__ rows) {
    __shared__ unsigned long long part[8];
    unsigned long long s = 0;
    for (uint32_t x = threadIdx.x; x < w; x += blockDim.x) s += counts[(uint64_t)blockIdx.x * w + x];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int q = 0; q < 8; q++) t += part[q];
        rows[blockIdx.x] = t;
    }
}

// =====================================================================================================
// microbenchmarks: what the FP32 / MUFU pipes of this GPU sustain (roofline denominators)
// =====================================================================================================
template <int KIND>
__global__ void __launch_bounds__(256) k_ubench(float* sink, int iters) {
    float a[8], s = 1.0f + threadIdx.x * 1e-7f, c = 0.5f + blockIdx.x * 1e-7f;
#pragma unroll
    for (int j = 0; j < 8; j++) a[j] = threadIdx.x * 0.001f + j;
    if (KIND == 0) {
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int j = 0; j < 8; j++) a[j] = fmaf(a[j], s, c);
        }
    } else if (KIND == 1) {
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int j = 0; j < 8; j++) a[j] = rcp_approx(a[j]) + c;
        }
    } else if (KIND == 3) {
        float2 b[8];
        const float2 s2 = make_float2(s, s), c2 = make_float2(c, c);
#pragma unroll
        for (int j = 0; j < 8; j++) b[j] = make_float2(a[j], a[j] + 0.5f);
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int j = 0; j < 8; j++) b[j] = __ffma2_rn(b[j], s2, c2);
        }
#pragma unroll
        for (int j = 0; j < 8; j++) a[j] = b[j].x + b[j].y;
    } else if (KIND == 5 || KIND == 6) {
        // FFMA2 and scalar FFMA interleaved (KIND 5: 8 + 8 per round, KIND 6: 8 + 4): do the two FMA sub-pipes take them
        // concurrently? counted as 32 "ops" per iteration like the others; the caller scales by the real flop count
        float2 b[8];
        const float2 s2 = make_float2(s, s), c2 = make_float2(c, c);
#pragma unroll
        for (int j = 0; j < 8; j++) b[j] = make_float2(a[j], a[j] + 0.5f);
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    b[j] = __ffma2_rn(b[j], s2, c2);
                    if (KIND == 5 || (j & 1)) a[j] = fmaf(a[j], s, c);
                }
        }
#pragma unroll
        for (int j = 0; j < 8; j++) a[j] += b[j].x + b[j].y;
    } else if (KIND >= 7 && KIND <= 9) {
        // 8 FFMA2 per round plus 2 MUFU.RCP (7), 2 FMNMX3 (8) or both (9): what the neighbours cost the FMA pipe
        float2 b[8];
        const float2 s2 = make_float2(s, s), c2 = make_float2(c, c);
        float m = 3.0e38f, q0 = s, q1 = c;
#pragma unroll
        for (int j = 0; j < 8; j++) b[j] = make_float2(a[j], a[j] + 0.5f);
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int r = 0; r < 4; r++) {
#pragma unroll
                for (int j = 0; j < 8; j++) b[j] = __ffma2_rn(b[j], s2, c2);
                if (KIND != 8) q0 = rcp_approx(q0), q1 = rcp_approx(q1);
                if (KIND != 7) m = fmin3(m, b[0].x, b[1].y), m = fmin3(m, b[2].x, b[3].y);
            }
        }
#pragma unroll
        for (int j = 0; j < 8; j++) a[j] += b[j].x + b[j].y;
        a[0] += m + q0 + q1;
    } else if (KIND == 4) {
        float2 X = make_float2(s, s), Y = make_float2(c, c), Z = make_float2(s + c, s + c);
        const float li = 1.0f;
        float2 px[4], py[4], pz[4], ac[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            px[j] = make_float2(a[2 * j], a[2 * j + 1]), py[j] = make_float2(a[2 * j] * 2.f, a[2 * j + 1] * 2.f);
            pz[j] = make_float2(a[2 * j] * 3.f, a[2 * j + 1] * 3.f), ac[j] = make_float2(0.f, 0.f);
        }
        const float2 inc = make_float2(0.25f, 0.25f);
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int r = 0; r < 4; r++) {
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const float2 dx = __fadd2_rn(X, px[j]), dy = __fadd2_rn(Y, py[j]), dz = __fadd2_rn(Z, pz[j]);
                    float2 d2 = __fmul2_rn(dx, dx);
                    d2 = __ffma2_rn(dy, dy, d2);
                    d2 = __ffma2_rn(dz, dz, d2);
                    const float r0 = rcp_approx(d2.x), r1 = rcp_approx(d2.y);
                    ac[j].x = (d2.x < 0.0001f) ? ac[j].x : fmaf(li, r0, ac[j].x);
                    ac[j].y = (d2.y < 0.0001f) ? ac[j].y : fmaf(li, r1, ac[j].y);
                }
                X = __fadd2_rn(X, inc), Y = __fadd2_rn(Y, inc), Z = __fadd2_rn(Z, inc);
            }
        }
#pragma unroll
        for (int j = 0; j < 4; j++) a[2 * j] = ac[j].x, a[2 * j + 1] = ac[j].y;
    } else {
        float lx = s, ly = c, lz = s + c, li = 1.0f;
        float px[8], py[8], pz[8];
#pragma unroll
        for (int j = 0; j < 8; j++) px[j] = a[j], py[j] = a[j] * 2.0f, pz[j] = a[j] * 3.0f, a[j] = 0.0f;
        for (int i = 0; i < iters; i++) {
#pragma unroll
            for (int r = 0; r < 4; r++) {
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const float dx = lx - px[j], dy = ly - py[j], dz = lz - pz[j];
                    const float d2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
                    const float rr = rcp_approx(d2);
                    a[j] = (d2 < 0.0001f) ? a[j] : fmaf(li, rr, a[j]);
                }
                lx += 0.25f, ly -= 0.125f, lz += 0.5f;  // 3 extra FADD per 8 evaluations (stands in for the LDS)
            }
        }
    }
    float t = 0.0f;
#pragma unroll
    for (int j = 0; j < 8; j++) t += a[j];
    if (t == 123.456f) sink[0] = t;  // never true: keeps the loop alive
}

}  // namespace

#ifdef VR_CHECKED
__global__ void k_check_selftest() { VR_CHECK(threadIdx.x > 1000u, 31); }  // always violated: proves the reporting path
void vrk_check_selftest(cudaStream_t st) { k_check_selftest<<<1, 32, 0, st>>>(); }
#endif
unsigned int vrk_check_flags() {
#ifdef VR_CHECKED
    unsigned int h = 0, zero = 0;
    cudaMemcpyFromSymbol(&h, g_vr_check, sizeof(h));
    if (h) cudaMemcpyToSymbol(g_vr_check, &zero, sizeof(zero));
    return h;
#else
    return 0;
#endif
}
static unsigned long long g_launches = 0;
unsigned long long vrk_launch_count() { return g_launches; }
#define VR_COUNT(n) (g_launches += (n))

double vrk_ubench(cudaStream_t st, int kind, float* sink, int iters) {
    VR_COUNT(1);
    const int blocks = 148 * 8, threads = 256;
    if (kind == 0) k_ubench<0><<<blocks, threads, 0, st>>>(sink, iters);
    else if (kind == 1) k_ubench<1><<<blocks, threads, 0, st>>>(sink, iters);
    else if (kind == 3) k_ubench<3><<<blocks, threads, 0, st>>>(sink, iters);
    else if (kind == 4) k_ubench<4><<<blocks, threads, 0, st>>>(sink, iters);
    else if (kind == 5) k_ubench<5><<<blocks, threads, 0, st>>>(sink, iters);
    else if (kind == 6) k_ubench<6><<<blocks, threads, 0, st>>>(sink, iters);
    else if (kind == 7) k_ubench<7><<<blocks, threads, 0, st>>>(sink, iters);
    else if (kind == 8) k_ubench<8><<<blocks, threads, 0, st>>>(sink, iters);
    else if (kind == 9) k_ubench<9><<<blocks, threads, 0, st>>>(sink, iters);
    else k_ubench<2><<<blocks, threads, 0, st>>>(sink, iters);
    return (double)blocks * threads * (double)iters * 32.0;  // 4 x 8 ops per iteration per thread
}

// =====================================================================================================
// launchers
// =====================================================================================================
void vrk_lightgen(cudaStream_t st, const VrGridDev& g, const vr_ubo& u, int keep_origin, uint32_t nx, uint32_t ny, uint32_t nz,
                  uint32_t* counts, unsigned long long* offsets, unsigned long long* total, unsigned long long* scan_tmp,
                  float4* slab, vr_ray_light* lights, VrLightCounter* counter) {
    VR_COUNT(4);
    const uint64_t n = (uint64_t)nx * ny * nz;
    const unsigned blocks = (unsigned)((n + 63) / 64);
    k_lightgen<VR_LG_MODE_LOCAL><<<blocks, 64, 0, st>>>(g, u, keep_origin, nx, ny, nz, counts, slab, nullptr, nullptr, nullptr);
    vrk_scan_u32(st, counts, offsets, total, n, scan_tmp);
    k_lightgen_compact<<<(unsigned)((n * VR_LG_LOCAL + 255) / 256), 256, 0, st>>>(u, n, counts, offsets, slab, lights, counter);
    k_lightgen<VR_LG_MODE_OVERFLOW><<<blocks, 64, 0, st>>>(g, u, keep_origin, nx, ny, nz, counts, nullptr, offsets, lights, counter);
    k_lightgen_finish<<<1, 1, 0, st>>>(u, total, counter);
}
uint32_t vrk_lightgen_local_capacity() { return VR_LG_LOCAL; }
void vrk_scan_u32(cudaStream_t st, const uint32_t* in, unsigned long long* out, unsigned long long* total, uint64_t n,
                  unsigned long long* tmp) {
    VR_COUNT(3);
    const uint64_t nblocks = (n + 4095) / 4096;
    if (nblocks == 0) {
        cudaMemsetAsync(total, 0, sizeof(unsigned long long), st);
        return;
    }
    k_scan_blocks<<<(unsigned)nblocks, 1024, 0, st>>>(in, out, tmp, n);
    k_scan_tops<<<1, 1024, 0, st>>>(tmp, nblocks, total);
    k_scan_add<<<(unsigned)nblocks, 1024, 0, st>>>(out, tmp, n);
}
void vrk_prepare_simple(cudaStream_t st, const vr_ray_light* lights, const VrLightCounter* counter, int use_from,
                        VrEvalLight* out, float4* out2, uint32_t capacity, unsigned long long* neval) {
    VR_COUNT(1);
    k_prepare_simple<<<(capacity + 255) / 256 + (capacity == 0), 256, 0, st>>>(lights, counter, use_from, out, out2, capacity, neval);
}
void vrk_prepare_count_steps(cudaStream_t st, const vr_ray_light* lights, const VrLightCounter* counter, float lrs,
                             uint32_t* steps, uint32_t capacity) {
    VR_COUNT(1);
    if (!capacity) return;
    k_prepare_count_steps<<<(capacity + 255) / 256, 256, 0, st>>>(lights, counter, lrs, steps, capacity);
}
void vrk_prepare_expand(cudaStream_t st, const vr_ray_light* lights, const VrLightCounter* counter, float lrs,
                        const unsigned long long* offsets, VrEvalLight* out, float4* out2, uint32_t capacity,
                        uint64_t out_capacity, uint32_t* overflow) {
    VR_COUNT(1);
    if (!capacity) return;
    k_prepare_expand<<<capacity, 64, 0, st>>>(lights, counter, lrs, offsets, out, out2, capacity, out_capacity, overflow);
}
void vrk_march(cudaStream_t st, const VrGridDev& g, const vr_ubo& u, float tan_scale, vr_tile tile, int mode, int trilinear,
               uint32_t* counts, const unsigned long long* offsets, VrSample* samples, VrWorkCounters* wc) {
    VR_COUNT(1);
    dim3 grid((tile.w + 15) / 16, (tile.h + 7) / 8);
    const VrSegLists none{};
#define VR_MARCH(M, T) k_march<M, T><<<grid, 128, 0, st>>>(g, u, tan_scale, tile, counts, offsets, samples, wc, none, 0)
    if (trilinear) {
        if (mode == 0) VR_MARCH(0, true);
        else if (mode == 1) VR_MARCH(1, true);
        else VR_MARCH(2, true);
    } else {
        if (mode == 0) VR_MARCH(0, false);
        else if (mode == 1) VR_MARCH(1, false);
        else VR_MARCH(2, false);
    }
#undef VR_MARCH
}
void vrk_march_build(cudaStream_t st, const VrGridDev& g, const vr_ubo& u, float tan_scale, vr_tile tile, int trilinear,
                     uint32_t* counts, const unsigned long long* offsets, VrSample* samples, VrWorkCounters* wc, const VrSegLists& sl,
                     int pairs) {
    cudaMemsetAsync(sl.counts, 0, 8 * sizeof(unsigned long long), st);
    VR_COUNT(1);
    dim3 grid((tile.w + 15) / 16, (tile.h + 7) / 8);
    if (trilinear) k_march<2, true, true><<<grid, 128, 0, st>>>(g, u, tan_scale, tile, counts, offsets, samples, wc, sl, pairs);
    else k_march<2, false, true><<<grid, 128, 0, st>>>(g, u, tan_scale, tile, counts, offsets, samples, wc, sl, pairs);
}
void vrk_march_bound(cudaStream_t st, const VrGridDev& g, const vr_ubo& u, float tan_scale, vr_tile tile, uint32_t* bound) {
    VR_COUNT(1);
    dim3 grid((tile.w + 15) / 16, (tile.h + 7) / 8);
    k_march_bound<<<grid, 128, 0, st>>>(g, u, tan_scale, tile, bound);
}
// samples of a buffer with gaps -> compact list in pixel order (test hook vr_read_samples)
__global__ void __launch_bounds__(256) k_compact_samples(const uint32_t* __restrict__ counts, const unsigned long long* __restrict__ offsets,
                                                         const unsigned long long* __restrict__ compact_offsets, uint64_t npix,
                                                         const VrSample* __restrict__ samples, VrSample* __restrict__ out) {
    const uint64_t p = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (p >= npix) return;
    const float4* src = reinterpret_cast<const float4*>(samples) + offsets[p];
    float4* dst = reinterpret_cast<float4*>(out) + compact_offsets[p];
    for (uint32_t i = 0, n = counts[p]; i < n; i++) dst[i] = src[i];
}
void vrk_compact_samples(cudaStream_t st, const uint32_t* counts, const unsigned long long* offsets, const unsigned long long* compact_offsets,
                         uint64_t npix, const VrSample* samples, VrSample* out) {
    if (npix == 0) return;
    VR_COUNT(1);
    k_compact_samples<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(counts, offsets, compact_offsets, npix, samples, out);
}
// opt-in phase function / light->sample transmittance: one thread per sample, every rounding of the oracle kept
void vrk_eval_phys(cudaStream_t st, int sphere, const VrGridDev& g, const vr_ubo& u, const VrPhysics& phys, const VrSample* samples,
                   uint64_t nsamples, const VrEvalLight* lights, const unsigned long long* nlights_dev, float* out) {
    if (nsamples == 0) return;
    VR_COUNT(1);
    const unsigned grid = (unsigned)((nsamples + 127) / 128);
    if (sphere)
        k_eval_phys<true><<<grid, 128, 0, st>>>(g, u, phys, samples, nsamples, lights, nlights_dev, out);
    else
        k_eval_phys<false><<<grid, 128, 0, st>>>(g, u, phys, samples, nsamples, lights, nlights_dev, out);
}
void vrk_trilinear(cudaStream_t st, const VrGridDev& g, const float* xyz, uint64_t n, float* out) {
    if (!n) return;
    VR_COUNT(1);
    k_trilinear<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(g, xyz, n, out);
}
void vrk_eval_point_packed(cudaStream_t st, int variant, const VrSample* samples, uint64_t nsamples, const float4* lights2,
                           const unsigned long long* nlights_dev, float* out, unsigned long long* redo_count) {
    if (nsamples == 0) return;
    VR_COUNT(1);
    auto grid = [&](int S) { return (unsigned)((nsamples + (uint64_t)EVAL_THREADS * S - 1) / ((uint64_t)EVAL_THREADS * S)); };
    switch (variant) {
        case 2: k_eval_point3<8, 3, 2><<<grid(8), EVAL_THREADS, 0, st>>>(samples, nsamples, nullptr, nullptr, nullptr, lights2, nlights_dev, out, redo_count); break;
        case 3: k_eval_point3<6, 4, 2><<<grid(6), EVAL_THREADS, 0, st>>>(samples, nsamples, nullptr, nullptr, nullptr, lights2, nlights_dev, out, redo_count); break;
        default: k_eval_point2<8><<<grid(8), EVAL_THREADS, 0, st>>>(samples, nsamples, lights2, nlights_dev, out); break;
    }
}
void vrk_build_segments(cudaStream_t st, const uint32_t* counts, const unsigned long long* offsets, uint64_t npix, const VrSegLists& sl,
                        int pairs) {
    cudaMemsetAsync(sl.counts, 0, 8 * sizeof(unsigned long long), st);
    if (!npix) return;
    VR_COUNT(1);
    k_build_segments<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(counts, offsets, npix, sl, pairs);
}
// the exact kernels over `list[0 .. *end_dev)`; launched over `bound` entries, CTAs beyond the real length leave at once.
// Small launches use 4 samples per thread: twice the CTAs at half the duration. strided: a fixed small grid whose CTAs
// walk the list in strides (for the hand-over list, which is normally empty: a launch over its capacity would be
// thousands of CTAs that only read the length).
static void launch_exact_list(cudaStream_t st, int sphere, const VrSample* samples, uint64_t bound, bool small, bool strided,
                              const uint32_t* list, const unsigned long long* end_dev, const float4* lights2,
                              const unsigned long long* nlights_dev, float radius, float* out, unsigned long long* redo_count) {
    if (!bound) return;
    VR_COUNT(1);
    const uint64_t per_cta = (uint64_t)EVAL_THREADS * (small ? 4 : 8);
    unsigned grid = (unsigned)((bound + per_cta - 1) / per_cta);
    if (strided) {  // (small implied)
        if (grid > 296u) grid = 296u;
        if (sphere) k_eval_sphere2<4, true, true><<<grid, EVAL_THREADS, 0, st>>>(samples, 0, list, nullptr, end_dev, lights2, nlights_dev, radius, out);
        else k_eval_point3<4, 3, 2, true, true><<<grid, EVAL_THREADS, 0, st>>>(samples, 0, list, nullptr, end_dev, lights2, nlights_dev, out, redo_count);
        return;
    }
    if (sphere) {
        if (small)
            k_eval_sphere2<4, true><<<grid, EVAL_THREADS, 0, st>>>(samples, 0, list, nullptr, end_dev, lights2, nlights_dev, radius, out);
        else
            k_eval_sphere2<8, true><<<grid, EVAL_THREADS, 0, st>>>(samples, 0, list, nullptr, end_dev, lights2, nlights_dev, radius, out);
    } else if (small) {
        k_eval_point3<4, 3, 2, true><<<grid, EVAL_THREADS, 0, st>>>(samples, 0, list, nullptr, end_dev, lights2, nlights_dev, out, redo_count);
    } else {
        k_eval_point3<8, 3, 2, true><<<grid, EVAL_THREADS, 0, st>>>(samples, 0, list, nullptr, end_dev, lights2, nlights_dev, out, redo_count);
    }
}

// one launch of a segment kernel instance over a table of at most `cap` entries: all full units, or up to one resident
// wave of chunks in SEG_SPLIT slice units each; CTAs whose ticket is beyond the work leave at once (the table lengths
// live on the device). The resident wave ("slots") is asked from the occupancy calculator once per device, which is also
// where the kernel is opted into its dynamic shared memory (function attributes are per device: vr_group drives several).
// resident wave of the full-unit instance on the current device (cached per device)
template <int THREADS, int MINB, bool SPHERE, int LEN>
static uint32_t seg_wave() {
    static uint32_t cache[64];  // (benign race: every thread computes the same value)
    int dev = 0;
    cudaGetDevice(&dev);
    uint32_t slots = dev >= 0 && dev < 64 ? cache[dev] : 0;
    if (!slots) {
        auto full = k_eval_seg<THREADS, MINB, SPHERE, LEN, false>;
        auto slice = k_eval_seg<THREADS, MINB, SPHERE, LEN, true>;
        constexpr uint32_t smem_full = seg_smem_bytes(LEN, THREADS), smem_slice = seg_smem_bytes(0, THREADS);
        int sms = 0, per_sm = 0;
        cudaFuncSetAttribute(full, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_full);
        cudaFuncSetAttribute(slice, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_slice);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, full, THREADS, smem_full);
        slots = (uint32_t)(sms * (per_sm > 0 ? per_sm : 1));
        if (dev >= 0 && dev < 64) cache[dev] = slots;
    }
    return slots;
}
uint64_t vrk_seg_part_floats(int len) {
    const uint32_t slots = len == 16 ? seg_wave<256, 2, false, 16>() : max(seg_wave<256, 2, false, 8>(), seg_wave<256, 2, true, 8>());
    return (uint64_t)slots * SEG_SPLIT * (uint64_t)len * 512u;
}
template <int THREADS, int MINB, bool SPHERE, int LEN, bool SLICE, class... Args>
static void launch_seg(cudaStream_t st, uint64_t cap, Args... args) {
    if (!cap) return;
    auto kernel = k_eval_seg<THREADS, MINB, SPHERE, LEN, SLICE>;
    constexpr uint32_t smem = seg_smem_bytes(SLICE ? 0 : LEN, THREADS);
    const uint32_t slots = seg_wave<THREADS, MINB, SPHERE, LEN>();
    const uint64_t chunks = (cap + 2 * THREADS - 1) / (2 * THREADS);
    const uint64_t sliced = chunks < slots ? chunks : slots;
    const uint64_t grid = SLICE ? sliced * SEG_SPLIT : chunks - sliced;
    if (!grid) return;
    VR_COUNT(1);
    kernel<<<(unsigned)grid, THREADS, smem, st>>>(args..., slots);
}
// Evaluation pass of the ray-segment path. Main stream: near-light flags, the flagged entries onto the exact kernel's list,
// then the big segment kernel. Side stream (lower priority, forked once the list is complete): the odd 8-sample segments and
// the exact kernel over the list (leftovers + near-light segments) -- independent work whose CTAs fill the SMs that the big
// kernel's end leaves idle. After the join: the exact kernel on whatever the segment kernels handed over themselves
// (entries whose samples deviate too much from their line; normally none: a small grid-stride launch over the second list).
void vrk_eval_point_seg(cudaStream_t st, cudaStream_t side, cudaStream_t side2, cudaEvent_t ev_fork, cudaEvent_t ev_join, cudaEvent_t ev_join2,
                        const VrGridDev& g, const vr_ubo& u, float tan_scale, vr_tile tile, const VrSample* samples,
                        uint64_t nsamples, uint64_t npix, const uint32_t* counts, const unsigned long long* offsets, const VrSegLists& sl,
                        const VrEvalLight* lights, const float4* lights2, uint64_t nlights_capacity,
                        const unsigned long long* nlights_dev, float rho, int sphere, float radius, float* out,
                        unsigned long long* redo_count) {
    if (nsamples == 0) return;
    if (!side) side = st;
    if (!side2) side2 = side;
    const uint64_t seg_cap = nsamples / SEG_LEN;
    const bool small = nsamples < 40ull * 1000 * 1000;  // short lists: 4 samples per thread, twice the CTAs at half the duration
    const uint64_t cap16 = sphere ? 0 : seg_cap / 2;
    const uint64_t cap8 = sphere ? seg_cap : (seg_cap < npix ? seg_cap : npix);  // point-type: at most one odd segment per pixel
    cudaMemsetAsync(sl.sched16, 0, sl.sched_bytes, st);  // the give-up flag (word 0) and the per-chunk slice counters of both tables
    if (seg_cap) {
        cudaMemsetAsync(sl.flags, 0, seg_cap, st);
        // sphere-type lists: distances are measured from the sphere's surface, the flags from its centre
        const float margin = 1.02f * (rho + (sphere ? radius : 0.0f)) + 1.0e-3f;
        VR_COUNT(1);  // (also restarts the exact kernel's list: launched even for an empty light list)
        const unsigned nb = (unsigned)(nlights_capacity < 65536u ? (nlights_capacity ? nlights_capacity : 1) : 65536u);
        k_flag_near_segments<<<dim3(nb, 4), 128, 0, st>>>(lights, nlights_dev, u, tan_scale, tile, counts, offsets, samples, sl, margin);
        VR_COUNT(1);
        const unsigned b16 = (unsigned)((cap16 + 255) / 256), b8 = (unsigned)((cap8 + 255) / 256);
        k_list_flagged<<<b16 + b8, 256, 0, st>>>(sl, b16);
    } else {
        VR_COUNT(1);
        k_seg_begin<<<1, 1, 0, st>>>(sl);
    }
    if (side != st) {
        cudaEventRecord(ev_fork, st);
        cudaStreamWaitEvent(side, ev_fork, 0);
        if (side2 != side) cudaStreamWaitEvent(side2, ev_fork, 0);
    }
    // Launch order = the order in which the pieces should start: the full units (longest) on the main stream; on the
    // lower-priority streams the slice units of the big table's last wave, the odd 8-sample segments, the exact kernel's
    // short CTAs.
    if (seg_cap) {
        if (sphere) {
            launch_seg<256, 2, true, 8, false>(st, cap8, g, u, tan_scale, tile, samples, sl, lights2, nlights_dev, rho, radius, out);
            launch_seg<256, 2, true, 8, true>(side, cap8, g, u, tan_scale, tile, samples, sl, lights2, nlights_dev, rho, radius, out);
        } else {
            launch_seg<256, 2, false, 16, false>(st, cap16, g, u, tan_scale, tile, samples, sl, lights2, nlights_dev, rho, 0.0f, out);
            launch_seg<256, 2, false, 16, true>(side, cap16, g, u, tan_scale, tile, samples, sl, lights2, nlights_dev, rho, 0.0f, out);
            launch_seg<256, 2, false, 8, false>(side2, cap8, g, u, tan_scale, tile, samples, sl, lights2, nlights_dev, rho, 0.0f, out);
            launch_seg<256, 2, false, 8, true>(side2, cap8, g, u, tan_scale, tile, samples, sl, lights2, nlights_dev, rho, 0.0f, out);
        }
    }
    // the list is complete since k_list_flagged (the segment kernels hand over into the second list)
    launch_exact_list(side2, sphere, samples, nsamples, small, false, sl.list, sl.counts + 2, lights2, nlights_dev, radius, out, redo_count);
    if (side != st) {
        cudaEventRecord(ev_join, side);
        cudaStreamWaitEvent(st, ev_join, 0);
        if (side2 != side) {
            cudaEventRecord(ev_join2, side2);
            cudaStreamWaitEvent(st, ev_join2, 0);
        }
    }
    launch_exact_list(st, sphere, samples, seg_cap * SEG_LEN, true, true, sl.list2, sl.counts + 7, lights2, nlights_dev, radius, out, redo_count);
}
void vrk_eval_sphere_packed(cudaStream_t st, const VrSample* samples, uint64_t nsamples, const float4* lights2,
                            const unsigned long long* nlights_dev, float radius, float* out) {
    if (nsamples == 0) return;
    VR_COUNT(1);
    constexpr int S = 8;
    const unsigned grid = (unsigned)((nsamples + (uint64_t)EVAL_THREADS * S - 1) / ((uint64_t)EVAL_THREADS * S));
    k_eval_sphere2<S, false><<<grid, EVAL_THREADS, 0, st>>>(samples, nsamples, nullptr, nullptr, nullptr, lights2, nlights_dev, radius, out);
}
void vrk_eval(cudaStream_t st, int sphere, const VrSample* samples, uint64_t nsamples, const VrEvalLight* lights,
              const unsigned long long* nlights_dev, float radius, float* out) {
    if (nsamples == 0) return;
    VR_COUNT(1);
    constexpr int S = 8;
    const uint64_t per_cta = (uint64_t)EVAL_THREADS * S;
    const unsigned grid = (unsigned)((nsamples + per_cta - 1) / per_cta);
    if (sphere)
        k_eval<true, S><<<grid, EVAL_THREADS, 0, st>>>(samples, nsamples, lights, nlights_dev, radius, out);
    else
        k_eval<false, S><<<grid, EVAL_THREADS, 0, st>>>(samples, nsamples, lights, nlights_dev, radius, out);
}
void vrk_resolve(cudaStream_t st, const vr_ubo& u, vr_tile tile, const uint32_t* counts, const unsigned long long* offsets,
                 const float* contrib, const VrLightCounter* counter, const VrFrameColorOut& frame_color, float4* accum, uchar4* rgba8,
                 int sum_mode) {
    VR_COUNT(1);
    const uint64_t npix = (uint64_t)tile.w * tile.h;
    if (!npix) return;
    k_resolve<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(u, tile, counts, offsets, contrib, counter, frame_color, accum, rgba8, sum_mode);
}
void vrk_accumulate(cudaStream_t st, const vr_ubo& u, vr_tile tile, const float* frame_color, float4* accum, uchar4* rgba8, int sum_mode) {
    VR_COUNT(1);
    const uint64_t npix = (uint64_t)tile.w * tile.h;
    if (!npix) return;
    k_accumulate<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(u, tile, frame_color, accum, rgba8, sum_mode);
}
void vrk_finish_sum(cudaStream_t st, float4* accum, uchar4* rgba8, uint64_t npix, float total) {
    if (!npix) return;
    VR_COUNT(1);
    k_finish_sum<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(accum, rgba8, npix, total);
}
void vrk_add_accum(cudaStream_t st, float4* dst, const float4* src, uint64_t npix) {
    if (!npix) return;
    VR_COUNT(1);
    k_add_accum<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(dst, src, npix);
}
void vrk_peer_barrier(cudaStream_t st, const VrPeerFlags& f, uint32_t self, uint32_t n, uint32_t epoch) {
    VR_COUNT(1);
    k_peer_barrier<<<1, 32, 0, st>>>(f, self, n, epoch, 10ull * 1000 * 1000 * 1000);
}
void vrk_path(cudaStream_t st, const VrGridDev& g, const vr_ubo& u, float tan_scale, vr_tile tile, const VrFrameColorOut& frame_color,
              float4* accum, uchar4* rgba8, VrWorkCounters* wc, float* tab, uint32_t nmax, int sum_mode) {
    VR_COUNT(1);
    if (tab) {
        VR_COUNT(1);
        k_path_table<<<1, 32, 0, st>>>(g, u, nmax, tab);
    }
    dim3 grid((tile.w + 15) / 16, (tile.h + 7) / 8);
    k_path<<<grid, 128, 0, st>>>(g, u, tan_scale, tile, frame_color, accum, rgba8, wc, tab, nmax, sum_mode);
}
void vrk_cpu_test(cudaStream_t st, const VrGridDev& g, uint32_t w, uint32_t h, float aspect, float scale, float* density) {
    VR_COUNT(1);
    dim3 grid((w + 15) / 16, (h + 7) / 8);
    k_cpu_test<<<grid, 128, 0, st>>>(g, w, h, aspect, scale, density);
}
void vrk_lookup(cudaStream_t st, const VrGridDev& g, const int32_t* ijk, uint64_t n, float* out) {
    VR_COUNT(1);
    if (!n) return;
    const uint32_t per_thread = 8;
    const uint64_t threads = (n + per_thread - 1) / per_thread;
    k_lookup<<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(g, ijk, n, per_thread, out);
}
void vrk_clear_image(cudaStream_t st, float4* accum, uchar4* rgba8, float* frame_color, uint64_t npix) {
    VR_COUNT(1);
    if (!npix) return;
    k_clear<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(accum, rgba8, frame_color, npix);
}
void vrk_row_sums(cudaStream_t st, const uint32_t* counts, uint32_t w, uint32_t h, unsigned long long* rows) {
    VR_COUNT(1);
    if (!h) return;
    k_row_sums<<<h, 256, 0, st>>>(counts, w, rows);
}
