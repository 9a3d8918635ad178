// vr_oracle.cpp -- CPU oracle for the many-light participating-media hot path.
//
// TEST INFRASTRUCTURE ONLY. Nothing in the product library links or calls this file; it is loaded by tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs as the checker.
//
// It restates, line by line, the arithmetic of the reference's compute shaders and of CPU_test:
//   shaders/common_functions.h:1-40, 182-201       hash / randf / randomDir / intersectAABB / light kernels
//   shaders/common_nanovdb.h:8-17                   grid -> tree -> root handles, grid type
//   shaders/PNanoVDB_GLSL.h:1276-1308, 2109-2231    root tile search, upper/lower/leaf descent
//   shaders/PNanoVDB_GLSL.h:2423-2487               float map transforms
//   shaders/light_gen.comp:9-100 (+ point_light_gen.comp:89-96 variant without the origin reset)
//   shaders/{point,sphere,ray,beam}_compute_color.comp, shaders/path_compute_color.comp
//   CPU_test/main.cpp:25-43, 89-126
//
// Parity status: the reference ships NO golden vectors for radiance or light lists, and
// its GLSL cannot be executed here (no graphics stack). The restatement is pinned two ways instead:
//   (1) oracle/_ref: the reference's OWN shader sources (.comp + common_*.h) compiled as C++ through a GLSL type
//       shim, with lookups through the reference's vendored PNanoVDB.h (C mode). tests/test_oracle_vs_ref.py
//       requires bit-identical light lists and frame colours between this file and that build, and the
//       golden digests under tests/golden/ were generated from it (tools/make_golden.py).
//   (2) the lookup layer against nanovdb::ReadAccessor / tree().getValue on NanoVDB-built grids.
// GLSL built-ins (exp, sin, cos, acos, tan, normalize) have implementation-defined precision; both this file
// and the shim take them from volumerenderer_b200/csrc/vr_glsl_math.h, which is what makes bit-identity with
// the CUDA kernels possible.
//
// Build with -ffp-contract=off (see oracle/Makefile): every float op must round individually.

#include "vr_oracle.h"

#include <atomic>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

#include "../volumerenderer_b200/csrc/vr_glsl_math.h"

#if defined(__FP_FAST_FMAF) && !defined(VRO_ALLOW_FMA_TARGET)
// -march=native makes FMA available; contraction must still be off (checked by the Makefile flag).
#endif

namespace {

// ------------------------------------------------------------------ GLSL-like scalar/vector helpers
struct V3 {
    float x, y, z;
};
inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
inline V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(float s, V3 a) { return V3{s * a.x, s * a.y, s * a.z}; }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float length(V3 a) { return vr_sqrt(dot(a, a)); }
inline V3 normalize(V3 a) {
    float l = length(a);
    return V3{a.x / l, a.y / l, a.z / l};
}
inline float gmin(float x, float y) { return (y < x) ? y : x; }  // GLSL min
inline float gmax(float x, float y) { return (x < y) ? y : x; }  // GLSL max
const float M_PI_F = (float)3.14159265358979323846;              // common_structures.h:2 narrowed to FP32

// ------------------------------------------------------------------ RNG  (common_functions.h:2-19)
inline uint32_t hash3(uint32_t vx, uint32_t vy, uint32_t vz) {
    uint32_t x = vx * 73856093u;
    uint32_t y = vy * 19349663u;
    uint32_t z = vz * 83492791u;
    uint32_t h = x ^ y ^ z;
    h = (h ^ (h >> 16u)) * 0x45d9f3bu;
    h = (h ^ (h >> 16u)) * 0x45d9f3bu;
    h = h ^ (h >> 16u);
    return h;
}
struct Seed {
    uint32_t x, y, z;
};
// float(hash) * (1.0/4294967295.0): the constant folds to 2^-32 in FP32
inline float randf(Seed s) { return (float)hash3(s.x, s.y, s.z) * (float)(1.0 / 4294967295.0); }
inline float randf_inc(Seed& s) {
    s.x += 1u;
    s.y += 1u;
    s.z += 1u;
    return randf(s);
}
// common_functions.h:21-25
inline V3 randomDir(float r1, float r2) {
    float theta = vr_acos(1.0f - 2.0f * r1);
    float phi = 2.0f * M_PI_F * r2;
    return normalize(v3(vr_sin(theta) * vr_cos(phi), vr_sin(theta) * vr_sin(phi), vr_cos(theta)));
}
// common_functions.h:27-40
inline bool intersectAABB(V3 origin, V3 dir, V3 boxMin, V3 boxMax, float& tmin, float& tmax) {
    const float o[3] = {origin.x, origin.y, origin.z};
    const float d[3] = {dir.x, dir.y, dir.z};
    const float bmin[3] = {boxMin.x, boxMin.y, boxMin.z};
    const float bmax[3] = {boxMax.x, boxMax.y, boxMax.z};
    for (int i = 0; i < 3; i++) {
        float invD = 1.0f / d[i];
        float t0 = (bmin[i] - o[i]) * invD;
        float t1 = (bmax[i] - o[i]) * invD;
        if (invD < 0.0f) {
            float temp = t0;
            t0 = t1;
            t1 = temp;
        }
        tmin = gmax(tmin, t0);
        tmax = gmin(tmax, t1);
        if (tmax < tmin) return false;
    }
    return true;
}

// ------------------------------------------------------------------ PNanoVDB float grid (PNanoVDB_GLSL.h)
// constants of the FLOAT grid type, PNanoVDB_GLSL.h:1192 and SURVEY Appendix D
enum : uint64_t {
    GRID_SIZE = 672,
    GRID_OFF_MAP = 296,
    GRID_OFF_GRID_TYPE = 636,
    MAP_OFF_MATF = 0,
    MAP_OFF_INVMATF = 36,
    MAP_OFF_VECF = 72,
    TREE_OFF_NODE_OFFSET_ROOT = 24,
    ROOT_OFF_BBOX_MIN = 0,
    ROOT_OFF_BBOX_MAX = 12,
    ROOT_OFF_TABLE_SIZE = 24,
    ROOT_OFF_BACKGROUND = 28,
    ROOT_SIZE = 64,
    ROOT_TILE_SIZE = 32,
    ROOT_TILE_OFF_KEY = 0,
    ROOT_TILE_OFF_CHILD = 8,
    ROOT_TILE_OFF_VALUE = 20,
    UPPER_OFF_CHILD_MASK = 4128,
    UPPER_OFF_TABLE = 8256,
    LOWER_OFF_CHILD_MASK = 544,
    LOWER_OFF_TABLE = 1088,
    LEAF_OFF_TABLE = 96,
};

struct Grid {
    const uint8_t* b;
    uint64_t tree, root;
    uint32_t gridType;
    int32_t bmin[3], bmax[3];
    float matf[9], invmatf[9], vecf[3];
};
inline uint32_t rd_u32(const uint8_t* b, uint64_t off) {
    uint32_t v;
    memcpy(&v, b + off, 4);
    return v;
}
inline int32_t rd_i32(const uint8_t* b, uint64_t off) { return (int32_t)rd_u32(b, off); }
inline uint64_t rd_u64(const uint8_t* b, uint64_t off) {
    uint64_t v;
    memcpy(&v, b + off, 8);
    return v;
}
inline float rd_f32(const uint8_t* b, uint64_t off) {
    float v;
    memcpy(&v, b + off, 4);
    return v;
}
// common_nanovdb.h:8-17 + PNanoVDB_GLSL.h:1239-1252
Grid open_grid(const void* buf) {
    Grid g;
    g.b = (const uint8_t*)buf;
    g.tree = GRID_SIZE;
    g.root = g.tree + rd_u64(g.b, g.tree + TREE_OFF_NODE_OFFSET_ROOT);
    g.gridType = rd_u32(g.b, GRID_OFF_GRID_TYPE);
    for (int i = 0; i < 3; i++) {
        g.bmin[i] = rd_i32(g.b, g.root + ROOT_OFF_BBOX_MIN + 4 * i);
        g.bmax[i] = rd_i32(g.b, g.root + ROOT_OFF_BBOX_MAX + 4 * i);
        g.vecf[i] = rd_f32(g.b, GRID_OFF_MAP + MAP_OFF_VECF + 4 * i);
    }
    for (int i = 0; i < 9; i++) {
        g.matf[i] = rd_f32(g.b, GRID_OFF_MAP + MAP_OFF_MATF + 4 * i);
        g.invmatf[i] = rd_f32(g.b, GRID_OFF_MAP + MAP_OFF_INVMATF + 4 * i);
    }
    return g;
}
// pnanovdb_root_find_tile + descent, PNanoVDB_GLSL.h:1276-1308, 2109-2194. The read accessor's cache only
// short-cuts this walk; the value returned is a pure function of ijk, so the oracle always walks from the root.
float lookup(const Grid& g, int32_t i, int32_t j, int32_t k) {
    const uint8_t* b = g.b;
    uint32_t tile_count = rd_u32(b, g.root + ROOT_OFF_TABLE_SIZE);
    uint64_t key = (uint64_t)((uint32_t)k >> 12) | ((uint64_t)((uint32_t)j >> 12) << 21) |
                   ((uint64_t)((uint32_t)i >> 12) << 42);
    uint64_t tile = g.root + ROOT_SIZE;
    bool found = false;
    for (uint32_t t = 0; t < tile_count; t++) {
        if (rd_u64(b, tile + ROOT_TILE_OFF_KEY) == key) {
            found = true;
            break;
        }
        tile += ROOT_TILE_SIZE;
    }
    if (!found) return rd_f32(b, g.root + ROOT_OFF_BACKGROUND);
    int64_t child = (int64_t)rd_u64(b, tile + ROOT_TILE_OFF_CHILD);
    if (child == 0) return rd_f32(b, tile + ROOT_TILE_OFF_VALUE);
    uint64_t upper = g.root + (uint64_t)child;
    uint32_t n = (((uint32_t)i & 4095u) >> 7) << 10 | (((uint32_t)j & 4095u) >> 7) << 5 | (((uint32_t)k & 4095u) >> 7);
    if (!((rd_u32(b, upper + UPPER_OFF_CHILD_MASK + 4u * (n >> 5)) >> (n & 31u)) & 1u))
        return rd_f32(b, upper + UPPER_OFF_TABLE + 8u * n);
    uint64_t lower = upper + rd_u64(b, upper + UPPER_OFF_TABLE + 8u * n);
    n = (((uint32_t)i & 127u) >> 3) << 8 | (((uint32_t)j & 127u) >> 3) << 4 | (((uint32_t)k & 127u) >> 3);
    if (!((rd_u32(b, lower + LOWER_OFF_CHILD_MASK + 4u * (n >> 5)) >> (n & 31u)) & 1u))
        return rd_f32(b, lower + LOWER_OFF_TABLE + 8u * n);
    uint64_t leaf = lower + rd_u64(b, lower + LOWER_OFF_TABLE + 8u * n);
    n = ((uint32_t)i & 7u) << 6 | ((uint32_t)j & 7u) << 3 | ((uint32_t)k & 7u);
    return rd_f32(b, leaf + LEAF_OFF_TABLE + 4u * n);
}
// PNanoVDB_GLSL.h:2436-2447 (apply_inverse), :2423-2434 (apply), :2461-2471 (apply_inverse_jacobi)
inline V3 world_to_index(const Grid& g, V3 s) {
    float sx = s.x - g.vecf[0], sy = s.y - g.vecf[1], sz = s.z - g.vecf[2];
    const float* m = g.invmatf;
    return v3(sx * m[0] + sy * m[1] + sz * m[2], sx * m[3] + sy * m[4] + sz * m[5], sx * m[6] + sy * m[7] + sz * m[8]);
}
inline V3 world_to_index_dir(const Grid& g, V3 s) {
    const float* m = g.invmatf;
    return v3(s.x * m[0] + s.y * m[1] + s.z * m[2], s.x * m[3] + s.y * m[4] + s.z * m[5],
              s.x * m[6] + s.y * m[7] + s.z * m[8]);
}
inline V3 index_to_world(const Grid& g, V3 s) {
    const float* m = g.matf;
    return v3(s.x * m[0] + s.y * m[1] + s.z * m[2] + g.vecf[0], s.x * m[3] + s.y * m[4] + s.z * m[5] + g.vecf[1],
              s.x * m[6] + s.y * m[7] + s.z * m[8] + g.vecf[2]);
}
inline bool in_bbox(const Grid& g, int32_t i, int32_t j, int32_t k) {
    return i >= g.bmin[0] && j >= g.bmin[1] && k >= g.bmin[2] && i <= g.bmax[0] && j <= g.bmax[1] && k <= g.bmax[2];
}

// ------------------------------------------------------------------ light kernels (common_functions.h:182-201)
inline float pointLight(V3 incomingPos, V3 lightPos, float lightIntensity) {
    V3 L = lightPos - incomingPos;
    float dist2 = dot(L, L);
    if (dist2 < 0.0001f) return 0.0f;
    return lightIntensity / (4.0f * M_PI_F * dist2);
}
inline V3 closestPointOnSphere(V3 sphereOrigin, float radius, V3 refPosition) {
    return sphereOrigin + normalize(refPosition - sphereOrigin) * radius;
}
inline float sphereLight(V3 incomingPos, V3 sphereOrigin, float lightIntensity, float beamRadius) {
    V3 curr = closestPointOnSphere(sphereOrigin, beamRadius, incomingPos);
    return pointLight(incomingPos, curr, lightIntensity);
}
inline V3 lfrom(const vro_light& l) { return v3(l.from[0], l.from[1], l.from[2]); }
inline V3 lto(const vro_light& l) { return v3(l.to[0], l.to[1], l.to[2]); }

// ------------------------------------------------------------------ opt-in physics
// None of this exists in the reference's shaders; the definitions below are this repository's own,
// stated here once and implemented identically by the CUDA kernels (csrc/vr_kernels.cu: vr_phys_factor, vr_trilinear).
//
// Trilinear density: nanovdb::math::TrilinearSampler (external/openvdb/nanovdb/nanovdb/math/SampleFromVoxels.h:174-260):
// ijk = floor(xyz), uvw = xyz - ijk, the eight values at ijk + {0,1}^3, interpolated along z, then y, then x with
// lerp(a, b, w) = a + w * (b - a)  (SampleFromVoxels.h:196-211, TrilinearSampler::sample).
inline float lerpf(float a, float b, float w) { return a + w * (b - a); }
float trilinear(const Grid& g, V3 xyz) {
    const float fx = vr_floor(xyz.x), fy = vr_floor(xyz.y), fz = vr_floor(xyz.z);
    const int32_t i = (int32_t)fx, j = (int32_t)fy, k = (int32_t)fz;
    const float u = xyz.x - fx, v = xyz.y - fy, w = xyz.z - fz;
    float c[2][2][2];
    for (int a = 0; a < 2; a++)
        for (int b = 0; b < 2; b++)
            for (int d = 0; d < 2; d++) c[a][b][d] = lookup(g, i + a, j + b, k + d);
    return lerpf(lerpf(lerpf(c[0][0][0], c[0][0][1], w), lerpf(c[0][1][0], c[0][1][1], w), v),
                 lerpf(lerpf(c[1][0][0], c[1][0][1], w), lerpf(c[1][1][0], c[1][1][1], w), v), u);
}
// Factor applied to the contribution of a light acting from `lightPos` on the sample at `p` (both world space):
//   phase:  Henyey-Greenstein relative to isotropic, P(c) = (1 - g^2) / (1 + g^2 - 2 g c)^(3/2), c = cos of the angle
//           between the light's travel direction (p - lightPos) and the direction to the camera (cam - p); P = 1 at g = 0
//   light -> sample transmittance: exp(-absorptionCoefficient * step * sum_{k<n} val(L + k*step*dir)) in index space, step =
//           rayMarchingStepSize, n = uint(distance / step), nearest-voxel values (the convention of the path mode's
//           shadow term, path_compute_color.comp:17-28)
float physFactor(const Grid& g, const vro_ubo& u, const vro_options& o, V3 p, V3 lightPos) {
    float f = 1.0f;
    if (o.phase_hg) {
        const V3 wi = normalize(p - lightPos);
        const V3 wo = normalize(v3(u.cameraPos[0], u.cameraPos[1], u.cameraPos[2]) - p);
        const float c = dot(wi, wo), gg = o.hg_g;
        const float den = (1.0f + gg * gg) - (2.0f * gg) * c;
        f *= (1.0f - gg * gg) / (den * vr_sqrt(den));
    }
    if (o.light_transmittance) {
        const V3 a = world_to_index(g, lightPos), b = world_to_index(g, p);
        const V3 ab = b - a;
        const float dist = length(ab);
        const float step = u.rayMarchingStepSize;
        const uint32_t n = (uint32_t)(dist / step);
        float sum = 0.0f;
        if (n > 0) {
            const V3 dir = ab * (1.0f / dist);
            for (uint32_t k = 0; k < n; k++) {
                const V3 q = a + dir * ((float)(int32_t)k * step);
                const int32_t i = (int32_t)vr_floor(q.x), j = (int32_t)vr_floor(q.y), kk = (int32_t)vr_floor(q.z);
                if (in_bbox(g, i, j, kk)) sum += lookup(g, i, j, kk);
            }
        }
        f *= vr_exp(-(u.absorptionCoefficient * step) * sum);
    }
    return f;
}

// {point,sphere,ray,beam}_compute_color.comp:9-26 computeLightContribution
float lightSum(int algo, const vro_ubo& u, const vro_light* lights, uint32_t lightCount, V3 p, uint64_t& evals,
               const Grid* g = nullptr, const vro_options* o = nullptr) {
    float total = 0.0f;
    if (o && (o->phase_hg || o->light_transmittance)) {  // same loops with the opt-in factor per evaluation
        for (uint32_t i = 0; i < lightCount; i++) {
            if (algo == VRO_POINT) {
                total += pointLight(p, lto(lights[i]), lights[i].intensity) * physFactor(*g, u, *o, p, lto(lights[i]));
                evals++;
            } else if (algo == VRO_SPHERE) {
                const V3 curr = closestPointOnSphere(lfrom(lights[i]), u.beamRadius, p);
                total += pointLight(p, curr, lights[i].intensity) * physFactor(*g, u, *o, p, curr);
                evals++;
            } else {
                V3 rayLocal = lto(lights[i]) - lfrom(lights[i]);
                V3 rayDir = normalize(rayLocal);
                uint32_t amountOfSteps = (uint32_t)(length(rayLocal) / u.lightRayStepSize);
                float pointIntensity = lights[i].intensity / (float)amountOfSteps;
                for (uint32_t s = 0; s < amountOfSteps; s++) {
                    V3 curr = lfrom(lights[i]) + ((float)(int32_t)s * u.lightRayStepSize) * rayDir;
                    if (algo == VRO_BEAM) curr = closestPointOnSphere(curr, u.beamRadius, p);
                    total += pointLight(p, curr, pointIntensity) * physFactor(*g, u, *o, p, curr);
                }
                evals += amountOfSteps;
            }
        }
        return total;
    }
    if (algo == VRO_POINT) {
        for (uint32_t i = 0; i < lightCount; i++) total += pointLight(p, lto(lights[i]), lights[i].intensity);
        evals += lightCount;
    } else if (algo == VRO_SPHERE) {
        // sphere_compute_color.comp:12 reads positionFrom
        for (uint32_t i = 0; i < lightCount; i++)
            total += sphereLight(p, lfrom(lights[i]), lights[i].intensity, u.beamRadius);
        evals += lightCount;
    } else {  // ray_compute_color.comp:11-25 / beam_compute_color.comp:11-25
        for (uint32_t i = 0; i < lightCount; i++) {
            V3 rayLocal = lto(lights[i]) - lfrom(lights[i]);
            V3 rayDir = normalize(rayLocal);
            uint32_t amountOfSteps = (uint32_t)(length(rayLocal) / u.lightRayStepSize);
            float pointIntensity = lights[i].intensity / (float)amountOfSteps;
            for (uint32_t s = 0; s < amountOfSteps; s++) {
                V3 curr = lfrom(lights[i]) + ((float)(int32_t)s * u.lightRayStepSize) * rayDir;
                if (algo == VRO_RAY)
                    total += pointLight(p, curr, pointIntensity);
                else
                    total += sphereLight(p, curr, pointIntensity, u.beamRadius);
            }
            evals += amountOfSteps;
        }
    }
    return total;
}

struct CamRay {
    V3 o, d;  // index space
    float tmin, tmax;
    bool hit;
};
// point_compute_color.comp:23-58 (identical in sphere/ray/beam/path)
CamRay cameraRay(const Grid& g, const vro_ubo& u, uint32_t px, uint32_t py) {
    uint32_t width = u.framebufferDim[0], height = u.framebufferDim[1];
    float aspectRatio = (float)width / (float)height;
    float scale = vr_tan(u.fov * 0.5f * M_PI_F / 180.0f);
    float x = (2.0f * ((float)(int32_t)px + 0.5f) / (float)width - 1.0f) * aspectRatio * scale;
    float y = (1.0f - 2.0f * ((float)(int32_t)py + 0.5f) / (float)height) * scale;
    V3 origin = v3(u.cameraPos[0], u.cameraPos[1], u.cameraPos[2]);
    V3 direction = normalize(v3(x, y, 1.0f));
    CamRay r;
    r.o = world_to_index(g, origin);
    r.d = normalize(world_to_index_dir(g, direction));
    V3 boxMin = v3((float)g.bmin[0], (float)g.bmin[1], (float)g.bmin[2]);
    V3 boxMax = v3((float)(g.bmax[0] + 1), (float)(g.bmax[1] + 1), (float)(g.bmax[2] + 1));
    r.tmin = 0.0f;
    r.tmax = u.rayMaxDistance;
    r.hit = intersectAABB(r.o, r.d, boxMin, boxMax, r.tmin, r.tmax);
    return r;
}

// point_compute_color.comp:60-97 -> frameColor for one pixel (all four light models)
float gatherPixel(const Grid& g, const vro_ubo& u, int algo, const vro_light* lights, uint32_t lightCount, uint32_t px,
                  uint32_t py, int skip_zero, vro_counters& c, const vro_options* o = nullptr) {
    CamRay r = cameraRay(g, u, px, py);
    float finalColor = 0.0f;
    if (r.hit) c.hit_pixels++;
    if (r.hit && r.tmax > 0.0f) {
        float tmin = gmax(r.tmin, 0.0f);
        float stepSize = u.rayMarchingStepSize;
        float t = tmin;
        float T = 1.0f;
        while (t < r.tmax && T > 0.001f) {
            V3 pLocal = r.o + r.d * t;
            V3 pWorld = index_to_world(g, pLocal);
            int32_t i = (int32_t)vr_floor(pLocal.x), j = (int32_t)vr_floor(pLocal.y), k = (int32_t)vr_floor(pLocal.z);
            if (in_bbox(g, i, j, k)) {
                float val = (o && o->trilinear) ? trilinear(g, pLocal) : lookup(g, i, j, k);
                c.S_ref++;
                if (val != 0.0f) c.S_eff++;
                if (!(skip_zero && val == 0.0f)) {
                    float lightIncoming = lightSum(algo, u, lights, lightCount, pWorld, c.E, &g, o);
                    float dC = T * val * stepSize * lightIncoming;
                    finalColor += dC;
                }
                float attenuation = vr_exp(-val * u.absorptionCoefficient * stepSize);
                T *= attenuation;
            }
            t += stepSize;
        }
    }
    float frameColor = finalColor / (float)lightCount;
    frameColor = gmin(gmax(frameColor, 0.0f), 1.0f);
    if (frameColor != frameColor) frameColor = 0.0f;  // NaN (lightCount == 0) stores as 0 (Appendix B)
    return frameColor;
}

// path_compute_color.comp:9-31
float pathLight(const Grid& g, const vro_ubo& u, V3 sampleLocalPos, float startIntensity) {
    float totalLight = startIntensity / 10000.0f;
    V3 ls = world_to_index(g, v3(u.lightSourceWorldPos[0], u.lightSourceWorldPos[1], u.lightSourceWorldPos[2]));
    V3 rayLocal = ls - sampleLocalPos;
    V3 rayDir = normalize(rayLocal);
    uint32_t amountOfSteps = (uint32_t)(length(rayLocal) / u.rayMarchingStepSize);
    for (uint32_t s = 0; s < amountOfSteps; s++) {
        V3 cur = ls + rayDir * u.rayMarchingStepSize;  // :20 -- the same position every iteration
        int32_t i = (int32_t)vr_floor(cur.x), j = (int32_t)vr_floor(cur.y), k = (int32_t)vr_floor(cur.z);
        float val = lookup(g, i, j, k);  // no bbox test (:22-24)
        float attenuation = vr_exp(-val * u.absorptionCoefficient * u.rayMarchingStepSize);
        totalLight *= attenuation;
    }
    return totalLight;
}
// path_compute_color.comp:33-117 -> frameColor
float pathPixel(const Grid& g, const vro_ubo& u, uint32_t px, uint32_t py, vro_counters& c) {
    Seed seed{px * u.frameCount, py * u.frameCount, 0u * u.frameCount};
    CamRay r = cameraRay(g, u, px, py);
    if (r.hit) c.hit_pixels++;
    float stepSize = u.rayMarchingStepSize;
    float T = 1.0f;
    float currentT = r.tmin;
    V3 o = r.o, d = r.d;
    V3 cur = o + d * r.tmin;
    float photonIntensity = u.photonInitialIntensity;
    float finalColor = 0.0f;
    while (r.hit && currentT < r.tmax && T > 0.001f && photonIntensity > 0.01f) {
        int32_t i = (int32_t)vr_floor(cur.x), j = (int32_t)vr_floor(cur.y), k = (int32_t)vr_floor(cur.z);
        if (in_bbox(g, i, j, k)) {
            float val = lookup(g, i, j, k);
            c.S_ref++;
            if (val > 0.0f) {
                c.S_eff++;
                float bounceRoll = randf_inc(seed);
                if (bounceRoll < u.scatteringProbability) {
                    float r1 = randf_inc(seed);
                    float r2 = randf_inc(seed);
                    d = randomDir(r1, r2);
                    o = cur;
                    currentT = 0.0f;
                }
                float lightIncoming = pathLight(g, u, cur, u.photonInitialIntensity);
                c.E++;
                float dC = T * val * stepSize * lightIncoming;
                finalColor += dC;
            }
        }
        currentT += stepSize;
        cur = o + d * currentT;
    }
    float frameColor = gmin(gmax(finalColor, 0.0f), 1.0f);
    if (frameColor != frameColor) frameColor = 0.0f;
    return frameColor;
}

template <class F>
void parallel_rows(uint32_t y0, uint32_t h, uint32_t stride, int nthreads, F&& f) {
    uint32_t nrows = (h + stride - 1) / stride;
    if (nthreads <= 1) {
        for (uint32_t r = 0; r < nrows; r++) f(y0 + r * stride, 0);
        return;
    }
    std::atomic<uint32_t> next{0};
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++)
        th.emplace_back([&, t]() {
            for (;;) {
                uint32_t r = next.fetch_add(1);
                if (r >= nrows) break;
                f(y0 + r * stride, t);
            }
        });
    for (auto& x : th) x.join();
}

}  // namespace

extern "C" {

uint32_t vro_hash(uint32_t x, uint32_t y, uint32_t z) { return hash3(x, y, z); }
float vro_randf(uint32_t x, uint32_t y, uint32_t z) { return randf(Seed{x, y, z}); }
void vro_random_dir(float r1, float r2, float out[3]) {
    V3 d = randomDir(r1, r2);
    out[0] = d.x;
    out[1] = d.y;
    out[2] = d.z;
}
float vro_exp(float x) { return vr_exp(x); }
float vro_sin(float x) { return vr_sin(x); }
float vro_cos(float x) { return vr_cos(x); }
float vro_acos(float x) { return vr_acos(x); }
float vro_tan(float x) { return vr_tan(x); }
// the same built-ins over an array (exhaustive sweeps in tests/test_builtins.py): which = 0 exp, 1 sin, 2 cos, 3 acos, 4 tan
int vro_builtin_array(int which, const float* in, uint64_t n, float* out) {
    if (which < 0 || which > 4) return 1;
    for (uint64_t i = 0; i < n; i++) {
        const float x = in[i];
        out[i] = which == 0 ? vr_exp(x) : which == 1 ? vr_sin(x) : which == 2 ? vr_cos(x) : which == 3 ? vr_acos(x) : vr_tan(x);
    }
    return 0;
}

int vro_grid_info(const void* grid, uint64_t size, int32_t bbox_min[3], int32_t bbox_max[3], float* background,
                  uint32_t node_counts[3], uint32_t* root_tiles) {
    if (!grid || size < GRID_SIZE + 64 + ROOT_SIZE) return 1;
    const uint8_t* b = (const uint8_t*)grid;
    uint64_t magic = rd_u64(b, 0);
    if (magic != 0x304244566f6e614eULL && magic != 0x314244566f6e614eULL && magic != 0x324244566f6e614eULL) return 2;
    if (rd_u32(b, GRID_OFF_GRID_TYPE) != 1u) return 3;
    Grid g = open_grid(grid);
    for (int i = 0; i < 3; i++) {
        bbox_min[i] = g.bmin[i];
        bbox_max[i] = g.bmax[i];
        if (node_counts) node_counts[i] = rd_u32(b, g.tree + 32 + 4 * i);
    }
    if (background) *background = rd_f32(b, g.root + ROOT_OFF_BACKGROUND);
    if (root_tiles) *root_tiles = rd_u32(b, g.root + ROOT_OFF_TABLE_SIZE);
    return 0;
}
void vro_trilinear(const void* grid, const float* xyz, uint64_t n, float* out) {
    Grid g = open_grid(grid);
    for (uint64_t i = 0; i < n; i++) out[i] = trilinear(g, v3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]));
}
void vro_lookup(const void* grid, const int32_t* ijk, uint64_t n, float* out) {
    Grid g = open_grid(grid);
    for (uint64_t q = 0; q < n; q++) out[q] = lookup(g, ijk[3 * q], ijk[3 * q + 1], ijk[3 * q + 2]);
}
void vro_world_to_index(const void* grid, const float* w, float* idx) {
    Grid g = open_grid(grid);
    V3 r = world_to_index(g, v3(w[0], w[1], w[2]));
    idx[0] = r.x;
    idx[1] = r.y;
    idx[2] = r.z;
}
void vro_index_to_world(const void* grid, const float* idx, float* w) {
    Grid g = open_grid(grid);
    V3 r = index_to_world(g, v3(idx[0], idx[1], idx[2]));
    w[0] = r.x;
    w[1] = r.y;
    w[2] = r.z;
}

// light_gen.comp:9-100; canonical order = photons by ((z*Ny)+y)*Nx+x, lights of a photon in emission order,
// list = first maxLights of the concatenation (one legal serialisation of the shader's atomicAdd race).
int vro_light_gen(const void* grid, const vro_ubo* ubo, int variant, uint32_t gx, uint32_t gy, uint32_t gz,
                  vro_light* lights_out, uint32_t* count_out, uint64_t* total_emitted, uint64_t* walk_steps) {
    Grid g = open_grid(grid);
    const vro_ubo& u = *ubo;
    uint32_t lightCount = 0;
    uint64_t emitted = 0, steps = 0;
    const uint32_t Nx = gx * 4u, Ny = gy * 4u, Nz = gz * 1u;  // local_size 4x4x1 (light_gen.comp:7)
    for (uint32_t z = 0; z < Nz; z++)
        for (uint32_t y = 0; y < Ny; y++)
            for (uint32_t x = 0; x < Nx; x++) {
                Seed seed{x * u.frameCount, y * u.frameCount, z * u.frameCount};
                V3 photonOriginWorld = v3(u.lightSourceWorldPos[0], u.lightSourceWorldPos[1], u.lightSourceWorldPos[2]);
                float r1 = randf_inc(seed);
                float r2 = randf_inc(seed);
                V3 photonDirWorld = randomDir(r1, r2);
                V3 o = world_to_index(g, photonOriginWorld);
                V3 d = normalize(world_to_index_dir(g, photonDirWorld));
                V3 boxMin = v3((float)g.bmin[0], (float)g.bmin[1], (float)g.bmin[2]);
                V3 boxMax = v3((float)(g.bmax[0] + 1), (float)(g.bmax[1] + 1), (float)(g.bmax[2] + 1));
                float tmin = 0.0f, tmax = u.rayMaxDistance;
                bool hitVolume = intersectAABB(o, d, boxMin, boxMax, tmin, tmax);
                float stepSize = u.rayMarchingStepSize;
                float T = 1.0f;
                float currentT = tmin;
                V3 cur = o + d * tmin;
                float photonIntensity = u.photonInitialIntensity;
                V3 previousPosition = photonOriginWorld;
                while (hitVolume && currentT < tmax && T > 0.001f && photonIntensity > 0.01f) {
                    steps++;
                    int32_t i = (int32_t)vr_floor(cur.x), j = (int32_t)vr_floor(cur.y), k = (int32_t)vr_floor(cur.z);
                    if (in_bbox(g, i, j, k)) {
                        float val = lookup(g, i, j, k);
                        if (val > 0.0f) {
                            float attenuation = vr_exp(-val * u.absorptionCoefficient * stepSize);
                            photonIntensity *= attenuation;
                            T *= attenuation;
                            float bounceRoll = randf_inc(seed);
                            if (bounceRoll < u.scatteringProbability) {
                                float q1 = randf_inc(seed);
                                float q2 = randf_inc(seed);
                                V3 newDirLocal = randomDir(q1, q2);
                                V3 w = index_to_world(g, cur);
                                uint32_t index = lightCount;  // atomicAdd(lightCount, 1)
                                lightCount += 1;
                                emitted++;
                                if (index < u.maxLights) {
                                    vro_light& L = lights_out[index];
                                    L.from[0] = previousPosition.x, L.from[1] = previousPosition.y, L.from[2] = previousPosition.z;
                                    L._pad0 = 0.0f;
                                    L.to[0] = w.x, L.to[1] = w.y, L.to[2] = w.z;
                                    L.intensity = photonIntensity;
                                    previousPosition = w;
                                } else {
                                    lightCount = u.maxLights;
                                }
                                d = newDirLocal;
                                if (variant == VRO_LG_RESET) {  // light_gen.comp:91-92
                                    o = cur;
                                    currentT = 0.0f;
                                }
                            }
                        }
                    }
                    currentT += stepSize;
                    cur = o + d * currentT;
                }
            }
    *count_out = lightCount;
    if (total_emitted) *total_emitted = emitted;
    if (walk_steps) *walk_steps = steps;
    return 0;
}

int vro_gather(const void* grid, const vro_ubo* ubo, int algo, const vro_light* lights, uint32_t light_count,
               uint32_t x0, uint32_t y0, uint32_t w, uint32_t h, uint32_t stride, int skip_zero, int nthreads,
               float* frame_color, vro_counters* counters) {
    return vro_gather_ex(grid, ubo, algo, lights, light_count, x0, y0, w, h, stride, skip_zero, nthreads, nullptr, frame_color,
                         counters);
}

int vro_gather_ex(const void* grid, const vro_ubo* ubo, int algo, const vro_light* lights, uint32_t light_count,
                  uint32_t x0, uint32_t y0, uint32_t w, uint32_t h, uint32_t stride, int skip_zero, int nthreads,
                  const vro_options* options, float* frame_color, vro_counters* counters) {
    Grid g = open_grid(grid);
    const vro_ubo u = *ubo;
    if (stride == 0) stride = 1;
    if (nthreads < 1) nthreads = 1;
    std::vector<vro_counters> cs((size_t)nthreads);
    for (auto& c : cs) memset(&c, 0, sizeof(c));
    const uint32_t W = u.framebufferDim[0];
    parallel_rows(y0, h, stride, nthreads, [&](uint32_t py, int tid) {
        vro_counters& c = cs[(size_t)tid];
        for (uint32_t px = x0; px < x0 + w; px += stride) {
            float fc = (algo == VRO_PATH) ? pathPixel(g, u, px, py, c)
                                          : gatherPixel(g, u, algo, lights, light_count, px, py, skip_zero, c, options);
            frame_color[(size_t)py * W + px] = fc;
            c.pixels++;
        }
    });
    if (counters) {
        memset(counters, 0, sizeof(*counters));
        for (auto& c : cs) {
            counters->S_ref += c.S_ref;
            counters->S_eff += c.S_eff;
            counters->E += c.E;
            counters->pixels += c.pixels;
            counters->hit_pixels += c.hit_pixels;
        }
    }
    return 0;
}

// point_compute_color.comp:99-105 (float n) and path_compute_color.comp:119-125 (uint n). rgba8: UNORM load k/255,
// store round-half-even(clamp(x,0,1)*255) (Appendix B).
void vro_accumulate(const float* frame_color, uint64_t n, uint32_t frame_count, int path_mode, float* accum_rgba,
                    uint8_t* rgba8) {
    float nm1, nn;
    if (path_mode) {
        nm1 = (float)(uint32_t)(frame_count - 1u);
        nn = (float)frame_count;
    } else {
        float totalInvocations = (float)(uint32_t)(frame_count * 1u);
        nm1 = totalInvocations - 1.0f;
        nn = totalInvocations;
    }
    for (uint64_t p = 0; p < n; p++) {
        float c = frame_color[p];
        if (accum_rgba) {
            for (int ch = 0; ch < 3; ch++) accum_rgba[4 * p + ch] = (accum_rgba[4 * p + ch] * nm1 + c) / nn;
            accum_rgba[4 * p + 3] = 1.0f;
        }
        if (rgba8) {
            for (int ch = 0; ch < 3; ch++) {
                float prev = (float)rgba8[4 * p + ch] / 255.0f;
                float v = (prev * nm1 + c) / nn;
                v = gmin(gmax(v, 0.0f), 1.0f);
                if (v != v) v = 0.0f;
                rgba8[4 * p + ch] = (uint8_t)nearbyintf(v * 255.0f);
            }
            rgba8[4 * p + 3] = 255;
        }
    }
}

// CPU_test/main.cpp:25-43 (raymarch) and :89-126 (pixel loop). glm::vec3 is replaced by three floats;
// glm::normalize(v) = v * (1/sqrt(dot(v,v))) (glm 0.9.9 func_geometric.inl; glm is not in the mount).
// tree().getValue(ijk) has no bbox clip: outside every root tile it returns the root background.
int vro_cpu_test(const void* grid, int width, int height, int nthreads, float* density_out, uint8_t* grey) {
    Grid g = open_grid(grid);
    const float camx = 0.0f, camy = 250.0f, camz = -800.0f;  // :92
    const float fov = 45.0f;                                 // :93
    const float aspectRatio = (float)width / height;         // :99
    const float scale = (float)tan(fov * 0.5 * M_PI / 180.0);  // :100 (double tan, narrowed)
    parallel_rows(0, (uint32_t)height, 1, nthreads, [&](uint32_t j, int) {
        for (int i = 0; i < width; ++i) {
            float x = (float)((2 * (i + 0.5) / width - 1) * aspectRatio * scale);  // :106 (double arithmetic)
            float y = (float)((1 - 2 * ((int)j + 0.5) / height) * scale);          // :107
            float inv = 1.0f / sqrtf(x * x + y * y + 1.0f * 1.0f);
            float dx = x * inv, dy = y * inv, dz = 1.0f * inv;
            const float tMax = 1200.0f, dt = 1.0f;  // :27-28
            float t = 0.0f, density = 0.0f;
            while (t < tMax) {
                float pxf = camx + t * dx, pyf = camy + t * dy, pzf = camz + t * dz;
                int32_t ii = (int32_t)floorf(pxf), jj = (int32_t)floorf(pyf), kk = (int32_t)floorf(pzf);
                float value = lookup(g, ii, jj, kk);
                density += value * dt;
                t += dt;
            }
            density_out[(size_t)j * width + i] = density;
            if (grey) {
                double m = density * 5.0;  // :119 std::min(density * 5.0, 255.0)
                if (m > 255.0) m = 255.0;
                grey[(size_t)j * width + i] = (uint8_t)m;
            }
        }
    });
    return 0;
}

}  // extern "C"
