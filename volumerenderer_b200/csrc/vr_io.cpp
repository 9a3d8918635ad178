// vr_io.cpp -- host-side file formats around the hot path. No CUDA here.
//
//   .nvdb input/output   the reference loads an OpenVDB file and converts it (src/main.cpp:1157-1193); OpenVDB is not
//                        available, so volumes come in as NanoVDB files: segments of FileHeader + FileMetaData + names +
//                        grids (external/openvdb/nanovdb/nanovdb/NanoVDB.h:6286-6325, io/IO.h:163-199, :381-437) or raw
//                        grid buffers (NanoVDB.h:6330-6334). Uncompressed codec only (ZIP/BLOSC need zlib/blosc).
//   PPM / PFM output     CPU_test writes a binary P6 file (CPU_test/main.cpp:128-132); the app itself only presents to a
//                        swapchain. PFM carries the linear float accumulator.
//   snapshots            progressive state = storage image + frameCount (src/main.cpp:657, :695-698, :937); saved and
//                        restored through the public read/write calls of vr.h.
#include <errno.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/vr.h"

namespace {

thread_local std::string g_io_error;

int io_fail(int code, const char* fmt, const char* a = "", unsigned long long b = 0) {
    char buf[600];
    snprintf(buf, sizeof(buf), fmt, a, b);
    g_io_error = buf;
    return code;
}

template <class T>
T rd(const uint8_t* p, uint64_t off) {
    T v;
    memcpy(&v, p + off, sizeof(T));
    return v;
}
template <class T>
void wr(uint8_t* p, uint64_t off, T v) {
    memcpy(p + off, &v, sizeof(T));
}

constexpr uint64_t kMagicNumb = 0x304244566f6e614eULL;  // "NanoVDB0": grids and files up to 32.6
constexpr uint64_t kMagicGrid = 0x314244566f6e614eULL;  // "NanoVDB1": raw grid buffers
constexpr uint64_t kMagicFile = 0x324244566f6e614eULL;  // "NanoVDB2": files
constexpr uint64_t kFileHeader = 16, kFileMeta = 176, kGridData = 672;

bool read_file(const char* path, std::vector<uint8_t>& out) {
    FILE* f = fopen(path, "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    const long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    if (n < 0) {
        fclose(f);
        return false;
    }
    out.resize((size_t)n);
    const size_t got = n ? fread(out.data(), 1, (size_t)n, f) : 0;
    fclose(f);
    return got == (size_t)n;
}

bool plausible_raw_grid(const uint8_t* p, uint64_t left) {
    if (left < kGridData + 64 + 64) return false;
    const uint64_t magic = rd<uint64_t>(p, 0);
    if (magic != kMagicNumb && magic != kMagicGrid) return false;
    const uint64_t size = rd<uint64_t>(p, 32);
    return size >= kGridData + 128 && size <= left && size % 32 == 0 && rd<uint32_t>(p, 636) < 64;
}

struct FoundGrid {
    uint64_t offset = 0, size = 0;
    std::string name;
    uint32_t type = 0;
};

// all grids of a file: segments (FileHeader, gridCount x (FileMetaData + name), then the grids) or raw buffers
int list_grids(const std::vector<uint8_t>& file, std::vector<FoundGrid>& out, const char* path) {
    const uint8_t* p = file.data();
    const uint64_t n = file.size();
    uint64_t pos = 0;
    while (pos + kFileHeader <= n) {
        const uint64_t magic = rd<uint64_t>(p, pos);
        bool as_file = (magic == kMagicFile);
        if (magic == kMagicNumb) {
            // ambiguous magic: a segment if its metadata is consistent with the file length, else a raw grid
            const uint16_t count = rd<uint16_t>(p, pos + 12);
            uint64_t q = pos + kFileHeader, total = 0;
            bool ok = count >= 1;
            for (uint16_t g = 0; ok && g < count; g++) {
                if (q + kFileMeta > n) {
                    ok = false;
                    break;
                }
                const uint64_t fsz = rd<uint64_t>(p, q + 8);
                const uint32_t name = rd<uint32_t>(p, q + 136);
                if (name > 4096 || fsz > n || total > n - fsz) {  // sums of 64-bit file fields must not wrap
                    ok = false;
                    break;
                }
                total += fsz;
                q += kFileMeta + name;
            }
            as_file = ok && q <= n && total <= n - q;
        }
        if (as_file) {
            const uint16_t count = rd<uint16_t>(p, pos + 12), codec = rd<uint16_t>(p, pos + 14);
            if (codec != 0)
                return io_fail(VR_E_BADGRID, "%s: compressed .nvdb (codec %llu: 1 = ZIP, 2 = BLOSC) is not supported; re-save it uncompressed", path, codec);
            uint64_t q = pos + kFileHeader;
            std::vector<FoundGrid> seg;
            for (uint16_t g = 0; g < count; g++) {
                if (q + kFileMeta > n) return io_fail(VR_E_BADGRID, "%s: truncated metadata", path);
                FoundGrid fg;
                fg.size = rd<uint64_t>(p, q + 0);
                const uint64_t fsz = rd<uint64_t>(p, q + 8);
                fg.type = rd<uint32_t>(p, q + 32);
                const uint32_t name = rd<uint32_t>(p, q + 136);
                if (rd<uint16_t>(p, q + 168) != 0 || fsz != fg.size)
                    return io_fail(VR_E_BADGRID, "%s: compressed grid inside the file is not supported", path);
                if (name > n || q + kFileMeta > n - name) return io_fail(VR_E_BADGRID, "%s: truncated grid name", path);
                if (name) fg.name.assign((const char*)p + q + kFileMeta, strnlen((const char*)p + q + kFileMeta, name));
                q += kFileMeta + name;
                seg.push_back(fg);
            }
            for (auto& fg : seg) {
                if (q > n || fg.size > n - q) return io_fail(VR_E_BADGRID, "%s: truncated grid data", path);
                fg.offset = q;
                q += fg.size;
                out.push_back(fg);
            }
            if (q <= pos) return io_fail(VR_E_BADGRID, "%s: malformed segment", path);  // the scan must advance
            pos = q;
        } else if (plausible_raw_grid(p + pos, n - pos)) {
            FoundGrid fg;
            fg.offset = pos, fg.size = rd<uint64_t>(p, pos + 32), fg.type = rd<uint32_t>(p, pos + 636);
            fg.name.assign((const char*)p + pos + 40, strnlen((const char*)p + pos + 40, 256));
            out.push_back(fg);
            pos += fg.size;  // plausible_raw_grid: kGridData + 128 <= size <= n - pos, so this advances and cannot wrap
        } else {
            if (out.empty()) return io_fail(VR_E_BADGRID, "%s: not a NanoVDB file or raw grid buffer", path);
            break;  // trailing bytes after the last grid
        }
    }
    if (out.empty()) return io_fail(VR_E_BADGRID, "%s: no grid found", path);
    return VR_OK;
}

}  // namespace

extern "C" {

const char* vr_io_last_error(void) { return g_io_error.c_str(); }

int vr_grid_load_nvdb(const char* path, const char* grid_name, void** bytes_out, uint64_t* size_out) {
    if (!path || !bytes_out || !size_out) return io_fail(VR_E_ARG, "vr_grid_load_nvdb: NULL argument");
    *bytes_out = nullptr, *size_out = 0;
    std::vector<uint8_t> file;
    if (!read_file(path, file)) return io_fail(VR_E_ARG, "cannot read %s", path);
    std::vector<FoundGrid> grids;
    int rc = list_grids(file, grids, path);
    if (rc != VR_OK) return rc;
    const FoundGrid* pick = nullptr;
    for (auto& g : grids) {
        if (g.type != 1u) continue;  // the first FloatGrid, like src/main.cpp:1173-1181
        if (grid_name && *grid_name && g.name != grid_name) continue;
        pick = &g;
        break;
    }
    if (!pick) return io_fail(VR_E_BADGRID, "%s: no float grid (GridType::Float) with the requested name among %llu grids", path, grids.size());
    // 32-byte aligned copy with gridIndex/gridCount rewritten for a single-grid buffer (io/IO.h:591 updateGridCount)
    void* mem = nullptr;
    if (posix_memalign(&mem, 32, pick->size) != 0) return io_fail(VR_E_CUDA, "out of host memory loading %s", path);
    memcpy(mem, file.data() + pick->offset, pick->size);
    wr<uint32_t>((uint8_t*)mem, 24, 0u);
    wr<uint32_t>((uint8_t*)mem, 28, 1u);
    *bytes_out = mem;
    *size_out = pick->size;
    return VR_OK;
}

int vr_grid_save_nvdb(const char* path, const void* bytes, uint64_t size, int raw) {
    if (!path || !bytes) return io_fail(VR_E_ARG, "vr_grid_save_nvdb: NULL argument");
    const uint8_t* g = (const uint8_t*)bytes;
    if (size < kGridData + 128 || !plausible_raw_grid(g, size)) return io_fail(VR_E_BADGRID, "vr_grid_save_nvdb: not a NanoVDB grid buffer");
    const uint64_t grid_size = rd<uint64_t>(g, 32);
    FILE* f = fopen(path, "wb");
    if (!f) return io_fail(VR_E_ARG, "cannot write %s", path);
    bool ok = true;
    if (!raw) {
        // one segment with one grid: FileHeader, FileMetaData, name, grid (NanoVDB.h:6352-6376)
        uint8_t head[kFileHeader] = {0}, meta[kFileMeta] = {0};
        const uint32_t version = rd<uint32_t>(g, 16);
        wr<uint64_t>(head, 0, kMagicNumb);
        wr<uint32_t>(head, 8, version);
        wr<uint16_t>(head, 12, 1);  // gridCount
        wr<uint16_t>(head, 14, 0);  // Codec::NONE
        const uint64_t tree = kGridData;
        const uint64_t root = tree + rd<uint64_t>(g, tree + 24);
        const char* name = (const char*)g + 40;
        const uint32_t name_size = (uint32_t)strnlen(name, 255) + 1;
        wr<uint64_t>(meta, 0, grid_size);                       // gridSize
        wr<uint64_t>(meta, 8, grid_size);                       // fileSize
        wr<uint64_t>(meta, 16, 0);                              // nameKey (0 = unchecked, io/IO.h:587)
        wr<uint64_t>(meta, 24, rd<uint64_t>(g, tree + 56));     // voxelCount
        wr<uint32_t>(meta, 32, rd<uint32_t>(g, 636));           // gridType
        wr<uint32_t>(meta, 36, rd<uint32_t>(g, 632));           // gridClass
        memcpy(meta + 40, g + 560, 48);                         // worldBBox
        memcpy(meta + 88, g + root, 24);                        // indexBBox = root bbox
        memcpy(meta + 112, g + 608, 24);                        // voxelSize
        wr<uint32_t>(meta, 136, name_size);
        for (int i = 0; i < 3; i++) wr<uint32_t>(meta, 140 + 4 * i, rd<uint32_t>(g, tree + 32 + 4 * i));  // nodeCount[leaf,lower,upper]
        wr<uint32_t>(meta, 152, 1u);                                                                       // one root
        for (int i = 0; i < 3; i++) wr<uint32_t>(meta, 156 + 4 * i, rd<uint32_t>(g, tree + 44 + 4 * i));  // tileCount
        wr<uint16_t>(meta, 168, 0);
        wr<uint16_t>(meta, 170, 0);
        wr<uint32_t>(meta, 172, version);
        std::string nm(name, name_size - 1);
        ok = fwrite(head, 1, sizeof(head), f) == sizeof(head) && fwrite(meta, 1, sizeof(meta), f) == sizeof(meta) &&
             fwrite(nm.c_str(), 1, name_size, f) == name_size;
    }
    ok = ok && fwrite(g, 1, grid_size, f) == grid_size;
    ok = (fclose(f) == 0) && ok;
    return ok ? VR_OK : io_fail(VR_E_ARG, "short write to %s", path);
}

int vr_image_write_ppm(const char* path, const uint8_t* rgba8, uint32_t width, uint32_t height) {
    if (!path || !rgba8 || !width || !height) return io_fail(VR_E_ARG, "vr_image_write_ppm: bad argument");
    FILE* f = fopen(path, "wb");
    if (!f) return io_fail(VR_E_ARG, "cannot write %s", path);
    fprintf(f, "P6\n%u %u\n255\n", width, height);  // CPU_test/main.cpp:130
    std::vector<uint8_t> row((size_t)width * 3);
    bool ok = true;
    for (uint32_t y = 0; y < height && ok; y++) {
        const uint8_t* src = rgba8 + (size_t)y * width * 4;
        for (uint32_t x = 0; x < width; x++) row[3 * x] = src[4 * x], row[3 * x + 1] = src[4 * x + 1], row[3 * x + 2] = src[4 * x + 2];
        ok = fwrite(row.data(), 1, row.size(), f) == row.size();
    }
    ok = (fclose(f) == 0) && ok;
    return ok ? VR_OK : io_fail(VR_E_ARG, "short write to %s", path);
}

int vr_image_write_pfm(const char* path, const float* data, uint32_t width, uint32_t height, uint32_t channels) {
    if (!path || !data || !width || !height || (channels != 1 && channels != 3 && channels != 4))
        return io_fail(VR_E_ARG, "vr_image_write_pfm: bad argument (channels must be 1, 3 or 4)");
    FILE* f = fopen(path, "wb");
    if (!f) return io_fail(VR_E_ARG, "cannot write %s", path);
    const uint32_t out_ch = channels == 1 ? 1 : 3;
    fprintf(f, "%s\n%u %u\n-1.0\n", out_ch == 1 ? "Pf" : "PF", width, height);  // negative scale = little endian
    std::vector<float> row((size_t)width * out_ch);
    bool ok = true;
    for (uint32_t yy = 0; yy < height && ok; yy++) {  // PFM rows run bottom to top
        const float* src = data + (size_t)(height - 1 - yy) * width * channels;
        for (uint32_t x = 0; x < width; x++)
            for (uint32_t c = 0; c < out_ch; c++) row[(size_t)x * out_ch + c] = src[(size_t)x * channels + c];
        ok = fwrite(row.data(), sizeof(float), row.size(), f) == row.size();
    }
    ok = (fclose(f) == 0) && ok;
    return ok ? VR_OK : io_fail(VR_E_ARG, "short write to %s", path);
}

// ---- snapshots: header + UBO + float4 accumulator + rgba8 image
struct VrSnapHeader {
    char magic[8];  // "VRSNAP01"
    uint32_t width, height, frame_count, reserved;
    uint64_t accum_bytes, rgba8_bytes;
    vr_ubo ubo;
};

int vr_snapshot_save(vr_ctx* ctx, const char* path) {
    if (!ctx || !path) return io_fail(VR_E_ARG, "vr_snapshot_save: NULL argument");
    vr_ubo u;
    int rc = vr_get_params(ctx, &u);
    if (rc != VR_OK) return rc;
    const uint32_t w = u.framebufferDim[0], h = u.framebufferDim[1];
    if (!w || !h) return io_fail(VR_E_STATE, "vr_snapshot_save: no image (call vr_resize first)");
    std::vector<float> accum((size_t)w * h * 4);
    std::vector<uint8_t> img((size_t)w * h * 4);
    rc = vr_read_accum_f32(ctx, accum.data());
    if (rc == VR_OK) rc = vr_read_image_rgba8(ctx, img.data());
    if (rc != VR_OK) return rc;
    VrSnapHeader hd{};
    memcpy(hd.magic, "VRSNAP01", 8);
    hd.width = w, hd.height = h, hd.frame_count = u.frameCount;
    hd.accum_bytes = accum.size() * sizeof(float), hd.rgba8_bytes = img.size();
    hd.ubo = u;
    FILE* f = fopen(path, "wb");
    if (!f) return io_fail(VR_E_ARG, "cannot write %s", path);
    bool ok = fwrite(&hd, sizeof(hd), 1, f) == 1 && fwrite(accum.data(), 1, hd.accum_bytes, f) == hd.accum_bytes &&
              fwrite(img.data(), 1, hd.rgba8_bytes, f) == hd.rgba8_bytes;
    ok = (fclose(f) == 0) && ok;
    return ok ? VR_OK : io_fail(VR_E_ARG, "short write to %s", path);
}

int vr_snapshot_load(vr_ctx* ctx, const char* path) {
    if (!ctx || !path) return io_fail(VR_E_ARG, "vr_snapshot_load: NULL argument");
    FILE* f = fopen(path, "rb");
    if (!f) return io_fail(VR_E_ARG, "cannot read %s", path);
    VrSnapHeader hd{};
    // the image limits of vr_resize apply before anything is sized from the header
    if (fread(&hd, sizeof(hd), 1, f) != 1 || memcmp(hd.magic, "VRSNAP01", 8) != 0 || hd.width == 0 || hd.height == 0 ||
        hd.width > 65536 || hd.height > 65536 || hd.accum_bytes != (uint64_t)hd.width * hd.height * 16 || hd.rgba8_bytes != (uint64_t)hd.width * hd.height * 4) {
        fclose(f);
        return io_fail(VR_E_ARG, "%s is not a libvr snapshot", path);
    }
    {   // the payload the header announces must be in the file before buffers are sized from it
        const long at = ftell(f);
        fseek(f, 0, SEEK_END);
        const long end = ftell(f);
        fseek(f, at, SEEK_SET);
        if (at < 0 || end < at || (uint64_t)(end - at) < hd.accum_bytes + hd.rgba8_bytes) {
            fclose(f);
            return io_fail(VR_E_ARG, "%s: truncated snapshot", path);
        }
    }
    std::vector<float> accum;
    std::vector<uint8_t> img;
    try {
        accum.resize((size_t)hd.width * hd.height * 4);
        img.resize((size_t)hd.width * hd.height * 4);
    } catch (...) {
        fclose(f);
        return io_fail(VR_E_ARG, "%s: out of host memory for the snapshot (%llu pixels)", path, (unsigned long long)hd.width * hd.height);
    }
    const bool ok = fread(accum.data(), 1, hd.accum_bytes, f) == hd.accum_bytes && fread(img.data(), 1, hd.rgba8_bytes, f) == hd.rgba8_bytes;
    fclose(f);
    if (!ok) return io_fail(VR_E_ARG, "%s: truncated snapshot", path);
    vr_ubo cur;
    int rc = vr_get_params(ctx, &cur);
    if (rc != VR_OK) return rc;
    if (cur.framebufferDim[0] != hd.width || cur.framebufferDim[1] != hd.height) {
        rc = vr_resize(ctx, hd.width, hd.height);
        if (rc != VR_OK) return rc;
    }
    rc = vr_set_params(ctx, &hd.ubo);
    if (rc == VR_OK) rc = vr_write_image_rgba8(ctx, img.data());
    if (rc == VR_OK) rc = vr_write_accum_f32(ctx, accum.data(), hd.frame_count);  // also restores frameCount
    return rc;
}

}  // extern "C"
