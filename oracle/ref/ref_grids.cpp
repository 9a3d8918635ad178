// ref_grids.cpp -- grid construction and value access through the reference's vendored NanoVDB (part of oracle/_ref).
//
// TEST INFRASTRUCTURE. Uses <reference>/external/openvdb/nanovdb headers where they lie (header-only, C++17):
//   * nanovdb::tools::createFogVolumeSphere / build::Grid functor + createNanoGrid -- the grid the reference's
//     createNanoGrid path would hand to the shaders (src/main.cpp:1191, CPU_test/main.cpp:85);
//   * NanoGrid<float>::tree().getValue and ReadAccessor -- the lookups CPU_test performs (CPU_test/main.cpp:37);
//   * the CPU_test ray-march loop (CPU_test/main.cpp:25-43, 89-126) restated around that getValue (the file itself
//     needs OpenVDB, glm and an MSVC header and cannot be compiled here).
#include <nanovdb/NanoVDB.h>
#include <nanovdb/tools/CreateNanoGrid.h>
#include <nanovdb/tools/CreatePrimitives.h>
#include <nanovdb/tools/GridBuilder.h>
#include <nanovdb/io/IO.h>
#include <nanovdb/math/SampleFromVoxels.h>

#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "../../volumerenderer_b200/csrc/vr_cloud.h"

namespace {
int copy_out(nanovdb::GridHandle<>& h, void** bytes, uint64_t* size) {
    if (!h) return 1;
    *size = h.size();
    *bytes = aligned_alloc(64, (h.size() + 63) & ~(size_t)63);
    memcpy(*bytes, h.data(), h.size());
    return 0;
}
}  // namespace

extern "C" {

void vrr_free(void* p) { free(p); }

// nanovdb::tools::createFogVolumeSphere<float>(radius, center, voxelSize, halfWidth, origin)
int vrr_make_fog_sphere(double radius, const double center[3], double voxel, double half_width, void** bytes, uint64_t* size) {
    auto h = nanovdb::tools::createFogVolumeSphere<float>(radius, nanovdb::Vec3d(center[0], center[1], center[2]), voxel,
                                                         half_width, nanovdb::Vec3d(0.0), "sphere_fog",
                                                         nanovdb::tools::StatsMode::Default, nanovdb::CheckMode::Disable);
    return copy_out(h, bytes, size);
}

// build::Grid<float>(0, FogVolume) filled by the vr_cloud functor over [0,N)^3, setTransform(voxel, translation)
int vrr_make_cloud(int32_t n, uint32_t seed, double voxel, const double tr[3], void** bytes, uint64_t* size) {
    nanovdb::tools::build::Grid<float> grid(0.0f, "cloud", nanovdb::GridClass::FogVolume);
    grid.setTransform(voxel, nanovdb::Vec3d(tr[0], tr[1], tr[2]));
    auto func = [=](const nanovdb::Coord& ijk) -> float { return vr_cloud_density(ijk[0], ijk[1], ijk[2], n, seed); };
    grid(func, nanovdb::CoordBBox(nanovdb::Coord(0), nanovdb::Coord(n - 1)));
    auto h = nanovdb::tools::createNanoGrid(grid, nanovdb::tools::StatsMode::Default, nanovdb::CheckMode::Disable);
    return copy_out(h, bytes, size);
}

// dense[(i*dims[1]+j)*dims[2]+k] at origin+ijk through the same functor route
int vrr_make_dense(const float* dense, const int32_t dims[3], const int32_t origin[3], float background, double voxel,
                   const double tr[3], void** bytes, uint64_t* size) {
    nanovdb::tools::build::Grid<float> grid(background, "dense", nanovdb::GridClass::FogVolume);
    grid.setTransform(voxel, nanovdb::Vec3d(tr[0], tr[1], tr[2]));
    const int32_t d0 = dims[0], d1 = dims[1], d2 = dims[2], o0 = origin[0], o1 = origin[1], o2 = origin[2];
    (void)d0;
    auto func = [=](const nanovdb::Coord& ijk) -> float {
        return dense[((size_t)(ijk[0] - o0) * d1 + (size_t)(ijk[1] - o1)) * d2 + (size_t)(ijk[2] - o2)];
    };
    grid(func, nanovdb::CoordBBox(nanovdb::Coord(o0, o1, o2), nanovdb::Coord(o0 + dims[0] - 1, o1 + dims[1] - 1, o2 + dims[2] - 1)));
    auto h = nanovdb::tools::createNanoGrid(grid, nanovdb::tools::StatsMode::Default, nanovdb::CheckMode::Disable);
    return copy_out(h, bytes, size);
}

// grid->tree().getValue(ijk) (mode 0, what CPU_test calls) or a ReadAccessor (mode 1) on ANY float grid buffer
int vrr_get_values(const void* grid_bytes, const int32_t* ijk, uint64_t n, int mode, float* out) {
    auto* grid = reinterpret_cast<const nanovdb::NanoGrid<float>*>(grid_bytes);
    if (!grid->isValid() || grid->gridType() != nanovdb::GridType::Float) return 1;
    if (mode == 0) {
        for (uint64_t q = 0; q < n; q++) out[q] = grid->tree().getValue(nanovdb::Coord(ijk[3 * q], ijk[3 * q + 1], ijk[3 * q + 2]));
    } else {
        auto acc = grid->getAccessor();
        for (uint64_t q = 0; q < n; q++) out[q] = acc.getValue(nanovdb::Coord(ijk[3 * q], ijk[3 * q + 1], ijk[3 * q + 2]));
    }
    return 0;
}

// nanovdb::math::createSampler<1>(accessor)(xyz): the vendored trilinear sampler (math/SampleFromVoxels.h:174-260, :366-375)
// at n index-space positions -- the CPU model of the opt-in trilinear density
int vrr_trilinear(const void* grid_bytes, const float* xyz, uint64_t n, float* out) {
    auto* grid = reinterpret_cast<const nanovdb::NanoGrid<float>*>(grid_bytes);
    if (!grid->isValid() || grid->gridType() != nanovdb::GridType::Float) return 1;
    auto acc = grid->getAccessor();
    auto smp = nanovdb::math::createSampler<1>(acc);
    for (uint64_t q = 0; q < n; q++) out[q] = smp(nanovdb::Vec3f(xyz[3 * q], xyz[3 * q + 1], xyz[3 * q + 2]));
    return 0;
}

int vrr_grid_facts(const void* grid_bytes, int32_t bbox[6], uint64_t* active, uint32_t counts[3], float* background) {
    auto* grid = reinterpret_cast<const nanovdb::NanoGrid<float>*>(grid_bytes);
    if (!grid->isValid()) return 1;
    auto bb = grid->indexBBox();
    for (int a = 0; a < 3; a++) bbox[a] = bb.min()[a], bbox[3 + a] = bb.max()[a];
    *active = grid->activeVoxelCount();
    counts[0] = grid->tree().nodeCount(0), counts[1] = grid->tree().nodeCount(1), counts[2] = grid->tree().nodeCount(2);
    *background = grid->tree().background();
    return 0;
}

// CPU_test/main.cpp:25-43 + :89-126 around nanovdb's own tree().getValue. glm::vec3 -> three floats,
// glm::normalize(v) = v * inversesqrt(dot(v,v)) with inversesqrt(x) = 1/sqrt(x) (glm's scalar definition).
int vrr_cpu_test(const void* grid_bytes, int width, int height, int nthreads, float* density_out) {
    auto* grid = reinterpret_cast<const nanovdb::NanoGrid<float>*>(grid_bytes);
    if (!grid->isValid()) return 1;
    const float cam[3] = {0.0f, 250.0f, -800.0f};
    const float fov = 45.0;
    const float aspectRatio = static_cast<float>(width) / height;
    const float scale = tan(fov * 0.5 * M_PI / 180.0);
    std::atomic<int> next{0};
    auto worker = [&]() {
        for (;;) {
            int j = next.fetch_add(1);
            if (j >= height) break;
            for (int i = 0; i < width; ++i) {
                float x = (2 * (i + 0.5) / width - 1) * aspectRatio * scale;
                float y = (1 - 2 * (j + 0.5) / height) * scale;
                float inv = 1.0f / std::sqrt(x * x + y * y + 1.0f * 1.0f);
                float dir[3] = {x * inv, y * inv, 1.0f * inv};
                const float tMax = 1200.0;
                const float dt = 1.0;
                float t = 0.0;
                float density = 0.0;
                while (t < tMax) {
                    float p[3] = {cam[0] + t * dir[0], cam[1] + t * dir[1], cam[2] + t * dir[2]};
                    int ii = (int)std::floor(p[0]), jj = (int)std::floor(p[1]), kk = (int)std::floor(p[2]);
                    float value = grid->tree().getValue(ii, jj, kk);
                    density += value * dt;
                    t += dt;
                }
                density_out[(size_t)j * width + i] = density;
            }
        }
    };
    if (nthreads < 1) nthreads = 1;
    std::vector<std::thread> th;
    for (int t = 1; t < nthreads; t++) th.emplace_back(worker);
    worker();
    for (auto& t : th) t.join();
    return 0;
}

// nanovdb::io::writeGrid (uncompressed) / readGrid: the .nvdb container the product's vr_grid_load_nvdb / vr_grid_save_nvdb
// must interoperate with (external/openvdb/nanovdb/nanovdb/io/IO.h:427-470, :560-600)
int vrr_write_nvdb(const void* grid_bytes, uint64_t size, const char* path) {
    try {
        auto buffer = nanovdb::HostBuffer::create(size);
        memcpy(buffer.data(), grid_bytes, size);
        nanovdb::GridHandle<> h(std::move(buffer));
        nanovdb::io::writeGrid(path, h, nanovdb::io::Codec::NONE);
        return 0;
    } catch (const std::exception& e) {
        fprintf(stderr, "vrr_write_nvdb: %s\n", e.what());
        return 1;
    }
}
int vrr_read_nvdb(const char* path, void** bytes, uint64_t* size) {
    try {
        auto h = nanovdb::io::readGrid(path, 0);
        if (!h.grid<float>()) return 2;
        return copy_out(h, bytes, size);
    } catch (const std::exception& e) {
        fprintf(stderr, "vrr_read_nvdb: %s\n", e.what());
        return 1;
    }
}

}  // extern "C"
