// vr_device.cuh -- device-side building blocks shared by the kernels: the PNanoVDB float-grid traversal, the map
// transforms, the shaders' counter-based RNG, the slab test and the camera ray. Everything that feeds a branch
// decision (voxel selection, scatter rolls, T cut-offs) is computed with individually rounded FP32 operations
// (vr_glsl_math.h) in the order the reference shaders write them, so results are bit-identical with the CPU checker.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/vr.h"
#include "vr_glsl_math.h"

// Grid facts parsed on the host at vr_set_grid (shaders/common_nanovdb.h:8-17 does this per thread) and passed to
// kernels by value. Offsets are 64-bit: grids beyond the 4 GiB GLSL limit (shaders/PNanoVDB_GLSL.h:28-30) work.
struct VrGridDev {
    const uint8_t* base;
    uint64_t root;        // byte offset of RootData
    uint32_t tile_count;  // root table size
    int32_t bmin[3], bmax[3];  // root bbox (inclusive), shaders/light_gen.comp:34-35
    float mat[9], inv[9], vec[3];  // Map::matF / invMatF / vecF at GridData+296 (PNanoVDB_GLSL.h:2423-2487)
};

// float-grid layout constants (shaders/PNanoVDB_GLSL.h:1192)
#define VR_ROOT_OFF_BACKGROUND 28u
#define VR_ROOT_SIZE 64u
#define VR_TILE_SIZE 32u
#define VR_TILE_OFF_CHILD 8u
#define VR_TILE_OFF_VALUE 20u
#define VR_UPPER_OFF_CHILD_MASK 4128u
#define VR_UPPER_OFF_TABLE 8256u
#define VR_LOWER_OFF_CHILD_MASK 544u
#define VR_LOWER_OFF_TABLE 1088u
#define VR_LEAF_OFF_TABLE 96u

__device__ __forceinline__ uint32_t vr_ld_u32(const uint8_t* p) { return __ldg(reinterpret_cast<const uint32_t*>(p)); }
__device__ __forceinline__ float vr_ld_f32(const uint8_t* p) { return __ldg(reinterpret_cast<const float*>(p)); }
__device__ __forceinline__ uint64_t vr_ld_u64(const uint8_t* p) {
    return __ldg(reinterpret_cast<const unsigned long long*>(p));
}

// Read accessor: caches the last leaf / lower / upper node like pnanovdb_readaccessor_t
// (shaders/PNanoVDB_GLSL.h:2053-2103). Offsets of 0 mean "not cached" (offset 0 is the grid header, never a node).
// Invariant: every cached node contains `key`.
struct VrAcc {
    int32_t kx, ky, kz;
    uint64_t leaf, lower, upper;
};
__device__ __forceinline__ void vr_acc_init(VrAcc& a) {
    a.kx = a.ky = a.kz = 0x7FFFFFFF;
    a.leaf = a.lower = a.upper = 0;
}

// Nearest-voxel value incl. tile and background values; the value mask is not consulted
// (pnanovdb_readaccessor_get_value_address + pnanovdb_read_float, shaders/PNanoVDB_GLSL.h:2202-2231, :365-368).
// log2dim receives the log2 edge length of the aligned cube that shares this value: 0 = a leaf voxel, 3 = a tile of a
// lower node (8^3 voxels), 7 = a tile of an upper node (128^3), 12 = a root tile (4096^3), -1 = background (no cube).
__device__ __forceinline__ float vr_lookup_dim(const VrGridDev& g, VrAcc& a, int32_t i, int32_t j, int32_t k, int& log2dim) {
    const uint8_t* b = g.base;
    const uint32_t dirty = (uint32_t)((i ^ a.kx) | (j ^ a.ky) | (k ^ a.kz));
    uint64_t leaf = (dirty & ~7u) ? 0 : a.leaf;
    if (!leaf) {
        a.leaf = 0;
        uint64_t lower = (dirty & ~127u) ? 0 : a.lower;
        if (!lower) {
            a.lower = 0;
            uint64_t upper = (dirty & ~4095u) ? 0 : a.upper;
            if (!upper) {
                a.upper = 0;
                // root tile search, PNanoVDB_GLSL.h:1276-1308
                const uint64_t key = (uint64_t)((uint32_t)k >> 12) | ((uint64_t)((uint32_t)j >> 12) << 21) |
                                     ((uint64_t)((uint32_t)i >> 12) << 42);
                uint64_t tile = g.root + VR_ROOT_SIZE;
                uint32_t t = 0;
                for (; t < g.tile_count; t++, tile += VR_TILE_SIZE)
                    if (vr_ld_u64(b + tile) == key) break;
                if (t == g.tile_count) return log2dim = -1, vr_ld_f32(b + g.root + VR_ROOT_OFF_BACKGROUND);
                const int64_t child = (int64_t)vr_ld_u64(b + tile + VR_TILE_OFF_CHILD);
                if (child == 0) return log2dim = 12, vr_ld_f32(b + tile + VR_TILE_OFF_VALUE);
                upper = g.root + (uint64_t)child;
                a.upper = upper;
                a.kx = i, a.ky = j, a.kz = k;
            }
            const uint32_t n = ((((uint32_t)i & 4095u) >> 7) << 10) | ((((uint32_t)j & 4095u) >> 7) << 5) |
                               (((uint32_t)k & 4095u) >> 7);
            const uint32_t m = vr_ld_u32(b + upper + VR_UPPER_OFF_CHILD_MASK + 4u * (n >> 5));
            if (!((m >> (n & 31u)) & 1u)) return log2dim = 7, vr_ld_f32(b + upper + VR_UPPER_OFF_TABLE + 8u * n);
            lower = upper + vr_ld_u64(b + upper + VR_UPPER_OFF_TABLE + 8u * n);
            a.lower = lower;
            a.kx = i, a.ky = j, a.kz = k;
        }
        const uint32_t n = ((((uint32_t)i & 127u) >> 3) << 8) | ((((uint32_t)j & 127u) >> 3) << 4) | (((uint32_t)k & 127u) >> 3);
        const uint32_t m = vr_ld_u32(b + lower + VR_LOWER_OFF_CHILD_MASK + 4u * (n >> 5));
        if (!((m >> (n & 31u)) & 1u)) return log2dim = 3, vr_ld_f32(b + lower + VR_LOWER_OFF_TABLE + 8u * n);
        leaf = lower + vr_ld_u64(b + lower + VR_LOWER_OFF_TABLE + 8u * n);
        a.leaf = leaf;
        a.kx = i, a.ky = j, a.kz = k;
    }
    const uint32_t n = (((uint32_t)i & 7u) << 6) | (((uint32_t)j & 7u) << 3) | ((uint32_t)k & 7u);
    log2dim = 0;
    return vr_ld_f32(b + leaf + VR_LEAF_OFF_TABLE + 4u * n);
}
__device__ __forceinline__ float vr_lookup(const VrGridDev& g, VrAcc& a, int32_t i, int32_t j, int32_t k) {
    int unused;
    return vr_lookup_dim(g, a, i, j, k, unused);
}

// ---- opt-in physics (not in the reference's shaders: definitions in oracle/vr_oracle.cpp and vr.h)
struct VrPhysics {
    uint32_t trilinear, phase_hg, light_transmittance;
    float hg_g;
};
// nanovdb::math::TrilinearSampler (external/openvdb/nanovdb/nanovdb/math/SampleFromVoxels.h:196-247): the eight values
// at floor(xyz) + {0,1}^3, lerp(a, b, w) = a + w * (b - a) along z, then y, then x
__device__ __forceinline__ float vr_lerp(float a, float b, float w) { return vr_add(a, vr_mul(w, vr_sub(b, a))); }
__device__ __forceinline__ float vr_trilinear(const VrGridDev& g, VrAcc& acc, vr_f3 xyz) {
    const float fx = vr_floor(xyz.x), fy = vr_floor(xyz.y), fz = vr_floor(xyz.z);
    const int32_t i = (int32_t)fx, j = (int32_t)fy, k = (int32_t)fz;
    const float u = vr_sub(xyz.x, fx), v = vr_sub(xyz.y, fy), w = vr_sub(xyz.z, fz);
    const float c000 = vr_lookup(g, acc, i, j, k), c001 = vr_lookup(g, acc, i, j, k + 1);
    const float c011 = vr_lookup(g, acc, i, j + 1, k + 1), c010 = vr_lookup(g, acc, i, j + 1, k);
    const float c100 = vr_lookup(g, acc, i + 1, j, k), c101 = vr_lookup(g, acc, i + 1, j, k + 1);
    const float c111 = vr_lookup(g, acc, i + 1, j + 1, k + 1), c110 = vr_lookup(g, acc, i + 1, j + 1, k);
    return vr_lerp(vr_lerp(vr_lerp(c000, c001, w), vr_lerp(c010, c011, w), v),
                   vr_lerp(vr_lerp(c100, c101, w), vr_lerp(c110, c111, w), v), u);
}

// Empty-space skipping that keeps the sample lattice. A camera ray at parameter t, position p = o + d*t (as the shader
// evaluates it in FP32), sits in a constant-value cube of edge 2^log2dim around voxel (i,j,k): a NanoVDB tile. Returns
// t_lim such that every later lattice point o + d*t', t <= t' < t_lim, floors into that same cube AND into the root
// bbox, whatever the FP32 rounding of o + d*t' does: the cube (clipped to the bbox) is shrunk by delta on every side,
// delta being orders of magnitude above the rounding error of a position (two roundings per axis at 1.2e-7 relative)
// and of t_lim itself. Positions move monotonically along each axis, so "inside the shrunk cube now" + "before the
// shrunk exit plane" covers the whole stretch. If p is within delta of a face, nothing is skipped (returns -inf).
// The caller still advances t by repeated `t += step` (the shader's own sequence), so every later sample position stays
// bit-identical; the skipped points are in-bbox samples of the tile's value.
__device__ __forceinline__ float vr_tile_exit(const VrGridDev& g, vr_f3 o, vr_f3 d, vr_f3 p, float t, int32_t i, int32_t j,
                                              int32_t k, int log2dim) {
    const int32_t mask = ~((1 << log2dim) - 1), dim = 1 << log2dim;
    const int32_t c[3] = {i, j, k};
    const float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z}, pp[3] = {p.x, p.y, p.z};
    float t_exit = 3.0e38f;
#pragma unroll
    for (int a = 0; a < 3; a++) {
        int32_t lo = c[a] & mask, hi = lo + dim;
        lo = lo > g.bmin[a] ? lo : g.bmin[a];
        hi = hi < g.bmax[a] + 1 ? hi : g.bmax[a] + 1;
        const float delta = 1.0e-3f + 1.0e-5f * (fabsf(pp[a]) + fabsf(oo[a])) + 2.0e-6f * fabsf(t);
        const float flo = (float)lo + delta, fhi = (float)hi - delta;
        if (!(pp[a] > flo && pp[a] < fhi)) return -3.0e38f;
        if (dd[a] > 0.0f) t_exit = fminf(t_exit, (fhi - oo[a]) / dd[a]);
        else if (dd[a] < 0.0f) t_exit = fminf(t_exit, (flo - oo[a]) / dd[a]);
    }
    return t_exit;
}

__device__ __forceinline__ bool vr_in_bbox(const VrGridDev& g, int32_t i, int32_t j, int32_t k) {
    return i >= g.bmin[0] && j >= g.bmin[1] && k >= g.bmin[2] && i <= g.bmax[0] && j <= g.bmax[1] && k <= g.bmax[2];
}

// pnanovdb_map_apply_inverse (PNanoVDB_GLSL.h:2436-2447): products summed left to right
__device__ __forceinline__ vr_f3 vr_world_to_index(const VrGridDev& g, vr_f3 s) {
    const float sx = vr_sub(s.x, g.vec[0]), sy = vr_sub(s.y, g.vec[1]), sz = vr_sub(s.z, g.vec[2]);
    const float* m = g.inv;
    return vr_make3(vr_add(vr_add(vr_mul(sx, m[0]), vr_mul(sy, m[1])), vr_mul(sz, m[2])),
                    vr_add(vr_add(vr_mul(sx, m[3]), vr_mul(sy, m[4])), vr_mul(sz, m[5])),
                    vr_add(vr_add(vr_mul(sx, m[6]), vr_mul(sy, m[7])), vr_mul(sz, m[8])));
}
// pnanovdb_map_apply_inverse_jacobi (PNanoVDB_GLSL.h:2461-2471)
__device__ __forceinline__ vr_f3 vr_world_to_index_dir(const VrGridDev& g, vr_f3 s) {
    const float* m = g.inv;
    return vr_make3(vr_add(vr_add(vr_mul(s.x, m[0]), vr_mul(s.y, m[1])), vr_mul(s.z, m[2])),
                    vr_add(vr_add(vr_mul(s.x, m[3]), vr_mul(s.y, m[4])), vr_mul(s.z, m[5])),
                    vr_add(vr_add(vr_mul(s.x, m[6]), vr_mul(s.y, m[7])), vr_mul(s.z, m[8])));
}
// pnanovdb_map_apply (PNanoVDB_GLSL.h:2423-2434)
__device__ __forceinline__ vr_f3 vr_index_to_world(const VrGridDev& g, vr_f3 s) {
    const float* m = g.mat;
    return vr_make3(vr_add(vr_add(vr_add(vr_mul(s.x, m[0]), vr_mul(s.y, m[1])), vr_mul(s.z, m[2])), g.vec[0]),
                    vr_add(vr_add(vr_add(vr_mul(s.x, m[3]), vr_mul(s.y, m[4])), vr_mul(s.z, m[5])), g.vec[1]),
                    vr_add(vr_add(vr_add(vr_mul(s.x, m[6]), vr_mul(s.y, m[7])), vr_mul(s.z, m[8])), g.vec[2]));
}

// ---- RNG, shaders/common_functions.h:2-19
struct VrSeed {
    uint32_t x, y, z;
};
__device__ __forceinline__ uint32_t vr_hash3(uint32_t vx, uint32_t vy, uint32_t vz) {
    uint32_t h = (vx * 73856093u) ^ (vy * 19349663u) ^ (vz * 83492791u);
    h = (h ^ (h >> 16)) * 0x45d9f3bu;
    h = (h ^ (h >> 16)) * 0x45d9f3bu;
    return h ^ (h >> 16);
}
__device__ __forceinline__ float vr_randf_inc(VrSeed& s) {
    s.x += 1u, s.y += 1u, s.z += 1u;
    return vr_mul(vr_u2f(vr_hash3(s.x, s.y, s.z)), 2.3283064365386963e-10f);  // float(h) * 2^-32 (Appendix B)
}
#define VR_PI_F 3.14159265358979323846f
// shaders/common_functions.h:21-25
__device__ __forceinline__ vr_f3 vr_random_dir(float r1, float r2) {
    const float theta = vr_acos(vr_sub(1.0f, vr_mul(2.0f, r1)));
    const float phi = vr_mul(vr_mul(2.0f, VR_PI_F), r2);
    const float st = vr_sin(theta), ct = vr_cos(theta), sp = vr_sin(phi), cp = vr_cos(phi);
    return vr_normalize3(vr_make3(vr_mul(st, cp), vr_mul(st, sp), ct));
}
// shaders/common_functions.h:27-40
__device__ __forceinline__ bool vr_intersect_aabb(vr_f3 o, vr_f3 d, const VrGridDev& g, float& tmin, float& tmax) {
    const float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
#pragma unroll
    for (int a = 0; a < 3; a++) {
        const float invD = vr_div(1.0f, dd[a]);
        float t0 = vr_mul(vr_sub(vr_i2f(g.bmin[a]), oo[a]), invD);
        float t1 = vr_mul(vr_sub(vr_i2f(g.bmax[a] + 1), oo[a]), invD);
        if (invD < 0.0f) {
            const float tmp = t0;
            t0 = t1;
            t1 = tmp;
        }
        tmin = vr_max(tmin, t0);
        tmax = vr_min(tmax, t1);
        if (tmax < tmin) return false;
    }
    return true;
}

// camera ray in index space, shaders/point_compute_color.comp:23-58 (same text in sphere/ray/beam/path)
// The reference camera has a position and a field of view only: it looks down +z, unrotated (:23-33), and the UBO has no
// field that could say otherwise. Code that inverts this projection (k_flag_near_segments maps a light's ball back to a
// pixel box) states its dependence on this constant, so that giving the camera an orientation cannot leave it behind.
#define VR_CAMERA_UNROTATED_PLUS_Z 1
struct VrCamRay {
    vr_f3 o, d;
    float tmin, tmax;
    bool hit;
};
__device__ __forceinline__ VrCamRay vr_camera_ray(const VrGridDev& g, const vr_ubo& u, float tan_scale, uint32_t px, uint32_t py) {
    const float width = vr_u2f(u.framebufferDim[0]), height = vr_u2f(u.framebufferDim[1]);
    const float aspect = vr_div(width, height);
    const float x = vr_mul(vr_mul(vr_sub(vr_div(vr_mul(2.0f, vr_add(vr_i2f((int32_t)px), 0.5f)), width), 1.0f), aspect), tan_scale);
    const float y = vr_mul(vr_sub(1.0f, vr_div(vr_mul(2.0f, vr_add(vr_i2f((int32_t)py), 0.5f)), height)), tan_scale);
    const vr_f3 dirW = vr_normalize3(vr_make3(x, y, 1.0f));
    VrCamRay r;
    r.o = vr_world_to_index(g, vr_make3(u.cameraPos[0], u.cameraPos[1], u.cameraPos[2]));
    r.d = vr_normalize3(vr_world_to_index_dir(g, dirW));
    r.tmin = 0.0f;
    r.tmax = u.rayMaxDistance;
    r.hit = vr_intersect_aabb(r.o, r.d, g, r.tmin, r.tmax);
    return r;
}
// tan(fov * 0.5 * M_PI / 180.0), identical on host and device
VR_HD float vr_fov_scale(float fov) { return vr_tan(vr_div(vr_mul(vr_mul(fov, 0.5f), 3.14159265358979323846f), 180.0f)); }

// Factor on the contribution of a light acting from lightPos on the sample at p (world space), oracle/vr_oracle.cpp
// physFactor: Henyey-Greenstein relative to isotropic for the angle between the light's travel direction and the
// direction to the camera, and exp(-sigma_a * step * sum of nearest-voxel values) over n = uint(distance / step) index-space
// steps from the light towards the sample. Individually rounded operations in the oracle's order.
__device__ __forceinline__ float vr_phys_factor(const VrGridDev& g, const vr_ubo& u, const VrPhysics& o, VrAcc& acc, vr_f3 p,
                                                vr_f3 lightPos) {
    float f = 1.0f;
    if (o.phase_hg) {
        const vr_f3 wi = vr_normalize3(vr_sub3(p, lightPos));
        const vr_f3 wo = vr_normalize3(vr_sub3(vr_make3(u.cameraPos[0], u.cameraPos[1], u.cameraPos[2]), p));
        const float c = vr_dot3(wi, wo), gg = o.hg_g;
        const float den = vr_sub(vr_add(1.0f, vr_mul(gg, gg)), vr_mul(vr_mul(2.0f, gg), c));
        f = vr_mul(f, vr_div(vr_sub(1.0f, vr_mul(gg, gg)), vr_mul(den, vr_sqrt(den))));
    }
    if (o.light_transmittance) {
        const vr_f3 a = vr_world_to_index(g, lightPos), b = vr_world_to_index(g, p);
        const vr_f3 ab = vr_sub3(b, a);
        const float dist = vr_length3(ab);
        const float step = u.rayMarchingStepSize;
        const uint32_t n = (uint32_t)vr_div(dist, step);
        float sum = 0.0f;
        if (n > 0) {
            const vr_f3 dir = vr_scale3(ab, vr_div(1.0f, dist));
            for (uint32_t k = 0; k < n; k++) {
                const vr_f3 q = vr_add3(a, vr_scale3(dir, vr_mul(vr_i2f((int32_t)k), step)));
                const int32_t i = (int32_t)vr_floor(q.x), j = (int32_t)vr_floor(q.y), kk = (int32_t)vr_floor(q.z);
                if (vr_in_bbox(g, i, j, kk)) sum = vr_add(sum, vr_lookup(g, acc, i, j, kk));
            }
        }
        f = vr_mul(f, vr_exp(vr_mul(-vr_mul(u.absorptionCoefficient, step), sum)));
    }
    return f;
}
