// vr_gridbuild.cpp -- host-side writer of NanoVDB float grids (raw buffer, the layout PNanoVDB reads).
//
// Replaces, for synthetic volumes, the openvdb -> nanovdb::tools::createNanoGrid conversion the reference performs at
// start-up (src/main.cpp:1157-1193, CPU_test/main.cpp:46-86). Written from the layout description in SURVEY.md
// Appendix D (structures: external/openvdb/nanovdb/nanovdb/NanoVDB.h:1844-1866 GridData, :2288-2294 TreeData,
// :2514-2594 RootData+Tile, :3198-3238 InternalData, :3773-3793 LeafData; mirrored by shaders/PNanoVDB_GLSL.h:1192).
// The result is value-identical (incl. root bbox and pruned uniform leaves) to what the vendored builder produces for
// the same input; tests/test_gridbuild.py checks that wherever the reference headers are available.
//
// Nodes are laid out breadth-first (root, upper, lower, leaf), 32-byte aligned; child links are byte offsets relative
// to the parent node (root tiles: relative to the root). Uniform leaves are pruned to lower-level tiles like
// tools::build::Grid::operator() does (tools/GridBuilder.h:1955-1985).
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/vr.h"
#include "vr_cloud.h"

namespace {

const uint64_t GRID_SIZE = 672, TREE_SIZE = 64, ROOT_SIZE = 64, TILE_SIZE = 32;
const uint64_t UPPER_SIZE = 270400, LOWER_SIZE = 33856, LEAF_SIZE = 2144;

struct LeafRec {
    int32_t ox, oy, oz;  // leaf origin (multiple of 8)
    int32_t uniform;     // 1: all 512 voxels equal `value` (!= background) -> lower-level tile
    float value;
    int64_t data;        // index into the owning thread's value store (units of 512 floats), -1 for uniform
    int32_t tid;
    uint64_t rootKey;
    uint32_t n2, n1;
};

inline uint64_t root_key(int32_t i, int32_t j, int32_t k) {
    return (uint64_t)((uint32_t)k >> 12) | ((uint64_t)((uint32_t)j >> 12) << 21) | ((uint64_t)((uint32_t)i >> 12) << 42);
}
inline uint32_t upper_slot(int32_t i, int32_t j, int32_t k) {
    return ((((uint32_t)i & 4095u) >> 7) << 10) | ((((uint32_t)j & 4095u) >> 7) << 5) | (((uint32_t)k & 4095u) >> 7);
}
inline uint32_t lower_slot(int32_t i, int32_t j, int32_t k) {
    return ((((uint32_t)i & 127u) >> 3) << 8) | ((((uint32_t)j & 127u) >> 3) << 4) | (((uint32_t)k & 127u) >> 3);
}

struct BBox {
    int32_t mn[3], mx[3];
    BBox() {
        for (int a = 0; a < 3; a++) mn[a] = INT32_MAX, mx[a] = INT32_MIN;
    }
    bool empty() const { return mn[0] > mx[0]; }
    void add(int32_t i, int32_t j, int32_t k) {
        const int32_t p[3] = {i, j, k};
        for (int a = 0; a < 3; a++) {
            if (p[a] < mn[a]) mn[a] = p[a];
            if (p[a] > mx[a]) mx[a] = p[a];
        }
    }
    void add(const BBox& b) {
        if (b.empty()) return;
        add(b.mn[0], b.mn[1], b.mn[2]);
        add(b.mx[0], b.mx[1], b.mx[2]);
    }
};
struct Stats {
    float mn = INFINITY, mx = -INFINITY;
    void add(float v) {
        if (v < mn) mn = v;
        if (v > mx) mx = v;
    }
    void add(const Stats& s) {
        add(s.mn);
        add(s.mx);
    }
};

template <class T>
inline void put(uint8_t* b, uint64_t off, T v) {
    memcpy(b + off, &v, sizeof(T));
}
inline void set_bit(uint8_t* mask, uint32_t n) { mask[n >> 3] |= (uint8_t)(1u << (n & 7u)); }

// Evaluates one 8^3 leaf: out[(di<<6)|(dj<<3)|dk] for the leaf with origin (ox,oy,oz).
struct LeafSource {
    int32_t lmin[3], lmax[3];  // inclusive range of leaf coordinates (origin >> 3)
    float background;
    virtual void eval(int32_t ox, int32_t oy, int32_t oz, float* out) const = 0;
    virtual ~LeafSource() {}
};

int build(const LeafSource& src, double voxel, const double tr[3], int nthreads, void** bytes_out, uint64_t* size_out) {
    if (!bytes_out || !size_out || !(voxel > 0.0)) return VR_E_ARG;
    if (nthreads < 1) nthreads = (int)std::thread::hardware_concurrency();
    if (nthreads < 1) nthreads = 1;
    const int64_t nx = (int64_t)src.lmax[0] - src.lmin[0] + 1, ny = (int64_t)src.lmax[1] - src.lmin[1] + 1,
                  nz = (int64_t)src.lmax[2] - src.lmin[2] + 1;
    if (nx <= 0 || ny <= 0 || nz <= 0) return VR_E_ARG;
    const int64_t total = nx * ny * nz;
    const float bg = src.background;

    // ---- pass 1: evaluate and classify leaves (parallel over slabs of leaves)
    std::vector<std::vector<LeafRec>> recs((size_t)nthreads);
    std::vector<std::vector<float>> store((size_t)nthreads);
    std::atomic<int64_t> next{0};
    const int64_t chunk = std::max<int64_t>(1, nz);
    auto worker = [&](int tid) {
        float vals[512];
        for (;;) {
            int64_t begin = next.fetch_add(chunk);
            if (begin >= total) break;
            int64_t end = std::min(total, begin + chunk);
            for (int64_t q = begin; q < end; q++) {
                int32_t lx = src.lmin[0] + (int32_t)(q / (ny * nz));
                int32_t ly = src.lmin[1] + (int32_t)((q / nz) % ny);
                int32_t lz = src.lmin[2] + (int32_t)(q % nz);
                int32_t ox = lx * 8, oy = ly * 8, oz = lz * 8;
                src.eval(ox, oy, oz, vals);
                bool allbg = true, uniform = true;
                for (int n = 0; n < 512; n++) {
                    if (vals[n] != bg) allbg = false;
                    if (vals[n] != vals[0]) uniform = false;
                }
                if (allbg) continue;
                LeafRec r;
                r.ox = ox, r.oy = oy, r.oz = oz;
                r.uniform = uniform ? 1 : 0;
                r.value = vals[0];
                r.tid = tid;
                r.rootKey = root_key(ox, oy, oz);
                r.n2 = upper_slot(ox, oy, oz);
                r.n1 = lower_slot(ox, oy, oz);
                if (uniform) {
                    r.data = -1;
                } else {
                    r.data = (int64_t)(store[(size_t)tid].size() / 512);
                    store[(size_t)tid].insert(store[(size_t)tid].end(), vals, vals + 512);
                }
                recs[(size_t)tid].push_back(r);
            }
        }
    };
    {
        std::vector<std::thread> th;
        for (int t = 1; t < nthreads; t++) th.emplace_back(worker, t);
        worker(0);
        for (auto& t : th) t.join();
    }
    std::vector<LeafRec> all;
    for (auto& v : recs) all.insert(all.end(), v.begin(), v.end());
    std::sort(all.begin(), all.end(), [](const LeafRec& a, const LeafRec& b) {
        if (a.rootKey != b.rootKey) return a.rootKey < b.rootKey;
        if (a.n2 != b.n2) return a.n2 < b.n2;
        return a.n1 < b.n1;
    });

    // ---- node census (breadth-first order = sorted order)
    uint64_t nUpper = 0, nLower = 0, nLeaf = 0;
    for (size_t q = 0; q < all.size(); q++) {
        bool newUpper = (q == 0) || all[q].rootKey != all[q - 1].rootKey;
        bool newLower = newUpper || all[q].n2 != all[q - 1].n2;
        nUpper += newUpper, nLower += newLower, nLeaf += all[q].uniform ? 0 : 1;
    }
    const uint64_t offTree = GRID_SIZE, offRoot = offTree + TREE_SIZE, offUpper = offRoot + ROOT_SIZE + TILE_SIZE * nUpper;
    const uint64_t offLower = offUpper + UPPER_SIZE * nUpper, offLeaf = offLower + LOWER_SIZE * nLower;
    const uint64_t size = offLeaf + LEAF_SIZE * nLeaf;
    uint8_t* b = (uint8_t*)aligned_alloc(64, (size + 63) & ~(uint64_t)63);
    if (!b) return VR_E_ARG;
    memset(b, 0, size);

    // ---- fill nodes
    BBox rootBox;
    Stats rootStats;
    uint64_t activeVoxels = 0, lowerTiles = 0;
    uint64_t iu = 0, il = 0, ileaf = 0;  // next free node of each level
    uint64_t curUpper = 0, curLower = 0;
    BBox upperBox, lowerBox;
    Stats upperStats, lowerStats;
    auto close_lower = [&]() {
        if (!curLower) return;
        for (int a = 0; a < 3; a++) put<int32_t>(b, curLower + 4 * a, lowerBox.mn[a]), put<int32_t>(b, curLower + 12 + 4 * a, lowerBox.mx[a]);
        put<float>(b, curLower + 1056, lowerStats.mn), put<float>(b, curLower + 1060, lowerStats.mx);
        upperBox.add(lowerBox), upperStats.add(lowerStats);
    };
    auto close_upper = [&]() {
        if (!curUpper) return;
        for (int a = 0; a < 3; a++) put<int32_t>(b, curUpper + 4 * a, upperBox.mn[a]), put<int32_t>(b, curUpper + 12 + 4 * a, upperBox.mx[a]);
        put<float>(b, curUpper + 8224, upperStats.mn), put<float>(b, curUpper + 8228, upperStats.mx);
        rootBox.add(upperBox), rootStats.add(upperStats);
    };
    for (size_t q = 0; q < all.size(); q++) {
        const LeafRec& r = all[q];
        bool newUpper = (q == 0) || r.rootKey != all[q - 1].rootKey;
        bool newLower = newUpper || r.n2 != all[q - 1].n2;
        if (newLower) close_lower();
        if (newUpper) {
            close_upper();
            curUpper = offUpper + UPPER_SIZE * iu;
            uint64_t tile = offRoot + ROOT_SIZE + TILE_SIZE * iu;
            put<uint64_t>(b, tile + 0, r.rootKey);
            put<int64_t>(b, tile + 8, (int64_t)(curUpper - offRoot));
            put<uint32_t>(b, tile + 16, 0u);
            put<float>(b, tile + 20, bg);
            // inactive slots of the tables hold the background value
            for (uint32_t n = 0; n < 32768; n++) put<float>(b, curUpper + 8256 + 8ull * n, bg);
            upperBox = BBox(), upperStats = Stats();
            iu++;
        }
        if (newLower) {
            curLower = offLower + LOWER_SIZE * il;
            put<int64_t>(b, curUpper + 8256 + 8ull * r.n2, (int64_t)(curLower - curUpper));
            set_bit(b + curUpper + 4128, r.n2);
            for (uint32_t n = 0; n < 4096; n++) put<float>(b, curLower + 1088 + 8ull * n, bg);
            lowerBox = BBox(), lowerStats = Stats();
            il++;
        }
        if (r.uniform) {  // active tile at the lower level
            put<uint64_t>(b, curLower + 1088 + 8ull * r.n1, 0ull);
            put<float>(b, curLower + 1088 + 8ull * r.n1, r.value);
            set_bit(b + curLower + 32, r.n1);
            lowerBox.add(r.ox, r.oy, r.oz), lowerBox.add(r.ox + 7, r.oy + 7, r.oz + 7);
            lowerStats.add(r.value);
            activeVoxels += 512, lowerTiles++;
            continue;
        }
        uint64_t leaf = offLeaf + LEAF_SIZE * ileaf++;
        put<int64_t>(b, curLower + 1088 + 8ull * r.n1, (int64_t)(leaf - curLower));
        set_bit(b + curLower + 544, r.n1);
        const float* vals = store[(size_t)r.tid].data() + 512 * (size_t)r.data;
        BBox lb;
        Stats ls;
        for (uint32_t n = 0; n < 512; n++) {
            if (vals[n] != bg) {
                set_bit(b + leaf + 16, n);
                lb.add(r.ox + (int32_t)(n >> 6), r.oy + (int32_t)((n >> 3) & 7u), r.oz + (int32_t)(n & 7u));
                ls.add(vals[n]);
                activeVoxels++;
            }
        }
        for (int a = 0; a < 3; a++) put<int32_t>(b, leaf + 4 * a, lb.mn[a]), put<uint8_t>(b, leaf + 12 + a, (uint8_t)(lb.mx[a] - lb.mn[a]));
        put<uint8_t>(b, leaf + 15, (uint8_t)((1u << 1) | (1u << 2)));  // has bbox | has min/max
        put<float>(b, leaf + 80, ls.mn), put<float>(b, leaf + 84, ls.mx);
        memcpy(b + leaf + 96, vals, 512 * sizeof(float));
        lowerBox.add(lb), lowerStats.add(ls);
    }
    close_lower();
    close_upper();

    // ---- root
    if (rootBox.empty()) {  // empty grid: NanoVDB stores an inverted bbox
        for (int a = 0; a < 3; a++) rootBox.mn[a] = INT32_MAX, rootBox.mx[a] = INT32_MIN;
        rootStats.mn = rootStats.mx = bg;
    }
    for (int a = 0; a < 3; a++) put<int32_t>(b, offRoot + 4 * a, rootBox.mn[a]), put<int32_t>(b, offRoot + 12 + 4 * a, rootBox.mx[a]);
    put<uint32_t>(b, offRoot + 24, (uint32_t)nUpper);
    put<float>(b, offRoot + 28, bg);
    put<float>(b, offRoot + 32, rootStats.mn), put<float>(b, offRoot + 36, rootStats.mx);
    // ---- tree
    put<int64_t>(b, offTree + 0, (int64_t)(offLeaf - offTree));
    put<int64_t>(b, offTree + 8, (int64_t)(offLower - offTree));
    put<int64_t>(b, offTree + 16, (int64_t)(offUpper - offTree));
    put<int64_t>(b, offTree + 24, (int64_t)(offRoot - offTree));
    put<uint32_t>(b, offTree + 32, (uint32_t)nLeaf), put<uint32_t>(b, offTree + 36, (uint32_t)nLower), put<uint32_t>(b, offTree + 40, (uint32_t)nUpper);
    put<uint32_t>(b, offTree + 44, (uint32_t)lowerTiles), put<uint32_t>(b, offTree + 48, 0u), put<uint32_t>(b, offTree + 52, 0u);
    put<uint64_t>(b, offTree + 56, activeVoxels);
    // ---- grid
    put<uint64_t>(b, 0, 0x304244566f6e614eULL);            // "NanoVDB0"
    put<uint64_t>(b, 8, ~0ull);                            // checksum disabled
    put<uint32_t>(b, 16, (32u << 21) | (7u << 10) | 0u);   // version 32.7.0
    put<uint32_t>(b, 20, (1u << 1) | (1u << 2) | (1u << 5)); // HasBBox | HasMinMax | IsBreadthFirst
    put<uint32_t>(b, 24, 0u), put<uint32_t>(b, 28, 1u);
    put<uint64_t>(b, 32, size);
    strncpy((char*)b + 40, "vr_synthetic", 255);
    const uint64_t map = 296;
    const float sF = (float)voxel, isF = (float)(1.0 / voxel);
    for (int d = 0; d < 3; d++) {
        put<float>(b, map + 0 + 4 * (4 * d), sF);
        put<float>(b, map + 36 + 4 * (4 * d), isF);
        put<float>(b, map + 72 + 4 * d, (float)tr[d]);
        put<double>(b, map + 88 + 8 * (4 * d), voxel);
        put<double>(b, map + 160 + 8 * (4 * d), 1.0 / voxel);
        put<double>(b, map + 232 + 8 * d, tr[d]);
    }
    put<float>(b, map + 84, 1.0f);
    put<double>(b, map + 256, 1.0);
    for (int d = 0; d < 3; d++) {
        double lo = rootBox.empty() ? 0.0 : (double)rootBox.mn[d] * voxel + tr[d];
        double hi = rootBox.empty() ? 0.0 : ((double)rootBox.mx[d] + 1.0) * voxel + tr[d];
        put<double>(b, 560 + 8 * d, lo), put<double>(b, 560 + 24 + 8 * d, hi);
        put<double>(b, 608 + 8 * d, voxel);
    }
    put<uint32_t>(b, 632, 2u);  // GridClass::FogVolume
    put<uint32_t>(b, 636, 1u);  // GridType::Float
    put<int64_t>(b, 640, (int64_t)size);
    put<uint32_t>(b, 648, 0u), put<uint32_t>(b, 652, 0u);
    put<uint64_t>(b, 656, 0ull);
    put<uint64_t>(b, 664, 0x314244566f6e614eULL);
    *bytes_out = b;
    *size_out = size;
    return VR_OK;
}

struct CloudSource : LeafSource {
    int32_t N;
    uint32_t seed;
    void eval(int32_t ox, int32_t oy, int32_t oz, float* out) const override {
        for (int di = 0; di < 8; di++)
            for (int dj = 0; dj < 8; dj++)
                for (int dk = 0; dk < 8; dk++) {
                    int32_t i = ox + di, j = oy + dj, k = oz + dk;
                    float v = (i < N && j < N && k < N) ? vr_cloud_density(i, j, k, N, seed) : 0.0f;
                    out[(di << 6) | (dj << 3) | dk] = v;
                }
    }
};
// fog sphere: 1 inside the core, linear ramp to 0 over a 3-voxel shell (not NanoVDB's createFogVolumeSphere profile;
// used for the CPU_test-style configuration where only the lookups matter)
struct SphereSource : LeafSource {
    double radius;
    int32_t c[3];
    void eval(int32_t ox, int32_t oy, int32_t oz, float* out) const override {
        for (int di = 0; di < 8; di++)
            for (int dj = 0; dj < 8; dj++)
                for (int dk = 0; dk < 8; dk++) {
                    float dx = (float)(ox + di - c[0]), dy = (float)(oy + dj - c[1]), dz = (float)(oz + dk - c[2]);
                    float d = sqrtf(dx * dx + dy * dy + dz * dz);
                    float v = ((float)radius - d) / 3.0f;
                    out[(di << 6) | (dj << 3) | dk] = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
                }
    }
};
struct DenseSource : LeafSource {
    const float* dense;
    int32_t dims[3], origin[3];
    void eval(int32_t ox, int32_t oy, int32_t oz, float* out) const override {
        for (int di = 0; di < 8; di++)
            for (int dj = 0; dj < 8; dj++)
                for (int dk = 0; dk < 8; dk++) {
                    int64_t i = (int64_t)ox + di - origin[0], j = (int64_t)oy + dj - origin[1], k = (int64_t)oz + dk - origin[2];
                    bool in = i >= 0 && j >= 0 && k >= 0 && i < dims[0] && j < dims[1] && k < dims[2];
                    out[(di << 6) | (dj << 3) | dk] = in ? dense[((size_t)i * dims[1] + (size_t)j) * dims[2] + (size_t)k] : background;
                }
    }
};

inline int32_t floordiv8(int32_t v) { return v >> 3; }

}  // namespace

extern "C" {

int vr_grid_build_cloud(int32_t n, uint32_t seed, double voxel_size, const double translation[3], int nthreads,
                        void** bytes_out, uint64_t* size_out) {
    if (n < 8 || n > 4096 || !translation) return VR_E_ARG;
    CloudSource s;
    s.N = n, s.seed = seed, s.background = 0.0f;
    for (int a = 0; a < 3; a++) s.lmin[a] = 0, s.lmax[a] = (n - 1) >> 3;
    return build(s, voxel_size, translation, nthreads, bytes_out, size_out);
}

int vr_grid_build_fog_sphere(double radius_vox, const int32_t center[3], double voxel_size, const double translation[3],
                             int nthreads, void** bytes_out, uint64_t* size_out) {
    if (!(radius_vox > 0.0) || radius_vox > 2000.0 || !center || !translation) return VR_E_ARG;
    SphereSource s;
    s.radius = radius_vox, s.background = 0.0f;
    int32_t r = (int32_t)ceil(radius_vox) + 1;
    for (int a = 0; a < 3; a++) s.c[a] = center[a], s.lmin[a] = floordiv8(center[a] - r), s.lmax[a] = floordiv8(center[a] + r);
    return build(s, voxel_size, translation, nthreads, bytes_out, size_out);
}

int vr_grid_build_dense(const float* dense, const int32_t dims[3], const int32_t origin[3], float background,
                        double voxel_size, const double translation[3], int nthreads, void** bytes_out,
                        uint64_t* size_out) {
    if (!dense || !dims || !origin || !translation) return VR_E_ARG;
    if (dims[0] < 1 || dims[1] < 1 || dims[2] < 1) return VR_E_ARG;
    DenseSource s;
    s.dense = dense, s.background = background;
    for (int a = 0; a < 3; a++) {
        s.dims[a] = dims[a], s.origin[a] = origin[a];
        s.lmin[a] = floordiv8(origin[a]), s.lmax[a] = floordiv8(origin[a] + dims[a] - 1);
    }
    return build(s, voxel_size, translation, nthreads, bytes_out, size_out);
}

void vr_grid_free(void* bytes) { free(bytes); }

}  // extern "C"
