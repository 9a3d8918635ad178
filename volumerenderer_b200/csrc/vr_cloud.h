// vr_cloud.h -- the synthetic procedural cloud of SURVEY.md section 8(d) / Appendix C (host only).
//
// density(i,j,k) = clamp(4*((1-r) + 1.6*(fbm-0.47) - 0.35), 0, 1)
//   r   = |ijk + 0.5 - N/2| / (N/2)
//   fbm = sum_{o<4} 0.5^(o+1) * valuenoise(ijk * (4/N) * 2^o, seed + o)
//   valuenoise: trilinear interpolation with smoothstep weights of lattice values hash(x,y,z,seed) * 2^-32
// Shared by the product's grid writer (vr_gridbuild.cpp) and by the test-only NanoVDB builder (oracle/ref) so the two
// can be compared voxel by voxel. Plain FP32 host arithmetic, contraction off.
#pragma once
#include <math.h>
#include <stdint.h>

static inline uint32_t vr_cloud_hash(int32_t x, int32_t y, int32_t z, uint32_t seed) {
    uint32_t h = (uint32_t)x * 73856093u ^ (uint32_t)y * 19349663u ^ (uint32_t)z * 83492791u ^ seed * 0x9E3779B9u;
    h = (h ^ (h >> 16)) * 0x45d9f3bu;
    h = (h ^ (h >> 16)) * 0x45d9f3bu;
    h = h ^ (h >> 16);
    return h;
}
static inline float vr_cloud_lattice(int32_t x, int32_t y, int32_t z, uint32_t seed) {
    return (float)vr_cloud_hash(x, y, z, seed) * 2.3283064365386963e-10f;
}
static inline float vr_cloud_smooth(float t) { return t * t * (3.0f - 2.0f * t); }
static inline float vr_cloud_noise(float x, float y, float z, uint32_t seed) {
    float fx = floorf(x), fy = floorf(y), fz = floorf(z);
    int32_t ix = (int32_t)fx, iy = (int32_t)fy, iz = (int32_t)fz;
    float tx = vr_cloud_smooth(x - fx), ty = vr_cloud_smooth(y - fy), tz = vr_cloud_smooth(z - fz);
    float c000 = vr_cloud_lattice(ix, iy, iz, seed), c100 = vr_cloud_lattice(ix + 1, iy, iz, seed);
    float c010 = vr_cloud_lattice(ix, iy + 1, iz, seed), c110 = vr_cloud_lattice(ix + 1, iy + 1, iz, seed);
    float c001 = vr_cloud_lattice(ix, iy, iz + 1, seed), c101 = vr_cloud_lattice(ix + 1, iy, iz + 1, seed);
    float c011 = vr_cloud_lattice(ix, iy + 1, iz + 1, seed), c111 = vr_cloud_lattice(ix + 1, iy + 1, iz + 1, seed);
    float x00 = c000 + (c100 - c000) * tx, x10 = c010 + (c110 - c010) * tx;
    float x01 = c001 + (c101 - c001) * tx, x11 = c011 + (c111 - c011) * tx;
    float y0 = x00 + (x10 - x00) * ty, y1 = x01 + (x11 - x01) * ty;
    return y0 + (y1 - y0) * tz;
}
static inline float vr_cloud_density(int32_t i, int32_t j, int32_t k, int32_t N, uint32_t seed) {
    const float half = 0.5f * (float)N;
    float dx = ((float)i + 0.5f - half) / half, dy = ((float)j + 0.5f - half) / half, dz = ((float)k + 0.5f - half) / half;
    float r = sqrtf(dx * dx + dy * dy + dz * dz);
    float fbm = 0.0f, amp = 0.5f, freq = 4.0f / (float)N;
    for (uint32_t o = 0; o < 4; o++) {
        fbm += amp * vr_cloud_noise((float)i * freq, (float)j * freq, (float)k * freq, seed + o);
        amp *= 0.5f;
        freq *= 2.0f;
    }
    float f = 4.0f * ((1.0f - r) + 1.6f * (fbm - 0.47f) - 0.35f);
    return f < 0.0f ? 0.0f : (f > 1.0f ? 1.0f : f);
}
