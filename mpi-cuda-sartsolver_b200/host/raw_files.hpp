// Flat little-endian stand-ins for the reference's HDF5 files, for machines without libhdf5 (this image has none):
// the same content per file -- one ray-transfer block per camera x voxel segment, one image series per camera, one
// Laplacian, one solution file -- without the container.
//
//   *.rtm  "SARTRTM1"  u64 pixel_offset, npixel, voxel_offset, nvoxel;  f32 value[npixel][nvoxel]
//          (source/raytransfer.cpp:92-110: dataset rtm/<name>/value of one camera x segment file)
//   *.img  "SARTIMG1"  u64 pixel_offset, npixel, nframes;  f64 time[nframes];  f64 frame[nframes][npixel]
//          (source/image.cpp:53-72, 286-305: image/time and the masked pixels of image/frame)
//   *.lap  "SARTLAP1"  u64 nvoxel, nnz;  u64 i[nnz];  u64 j[nnz];  f32 value[nnz]
//          (source/laplacian.cpp:40-65: laplacian/i, j, value)
//   out    "SARTSOL1"  u64 nvoxel, nframes;  per frame: f64 time, i32 status, i32 iterations, f64 value[nvoxel]
//          (source/solution.cpp:60-112: solution/value, time, status)
#pragma once

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace rawfiles {

constexpr char kRtmMagic[9] = "SARTRTM1";
constexpr char kImgMagic[9] = "SARTIMG1";
constexpr char kLapMagic[9] = "SARTLAP1";
constexpr char kSolMagic[9] = "SARTSOL1";

struct RtmHeader {
    std::string path;
    uint64_t pixel_offset = 0, npixel = 0, voxel_offset = 0, nvoxel = 0;
    static constexpr uint64_t kDataOffset = 8 + 4 * 8;
};

struct ImgHeader {
    std::string path;
    uint64_t pixel_offset = 0, npixel = 0, nframes = 0;
    std::vector<double> time;
    uint64_t data_offset() const { return 8 + 3 * 8 + nframes * 8; }
};

inline bool read_exact(FILE* f, void* dst, size_t bytes) { return bytes == 0 || std::fread(dst, 1, bytes, f) == bytes; }

// returns "" when the file cannot be read
inline std::string file_magic(const std::string& path) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return "";
    char m[9] = {0};
    const bool ok = read_exact(f, m, 8);
    std::fclose(f);
    return ok ? std::string(m) : "";
}

inline bool read_rtm_header(const std::string& path, RtmHeader& h) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    char m[9] = {0};
    uint64_t v[4];
    const bool ok = read_exact(f, m, 8) && std::strcmp(m, kRtmMagic) == 0 && read_exact(f, v, sizeof(v));
    std::fclose(f);
    if (ok) {
        h.path = path;
        h.pixel_offset = v[0];
        h.npixel = v[1];
        h.voxel_offset = v[2];
        h.nvoxel = v[3];
    }
    return ok;
}

inline bool read_img_header(const std::string& path, ImgHeader& h) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    char m[9] = {0};
    uint64_t v[3];
    bool ok = read_exact(f, m, 8) && std::strcmp(m, kImgMagic) == 0 && read_exact(f, v, sizeof(v));
    if (ok) {
        h.path = path;
        h.pixel_offset = v[0];
        h.npixel = v[1];
        h.nframes = v[2];
        h.time.resize(h.nframes);
        ok = read_exact(f, h.time.data(), h.nframes * 8);
    }
    std::fclose(f);
    return ok;
}

inline bool read_laplacian(const std::string& path, uint64_t& nvoxel, std::vector<uint64_t>& flat_index,
                           std::vector<float>& value) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    char m[9] = {0};
    uint64_t v[2];
    bool ok = read_exact(f, m, 8) && std::strcmp(m, kLapMagic) == 0 && read_exact(f, v, sizeof(v));
    if (ok) {
        nvoxel = v[0];
        std::vector<uint64_t> i(v[1]), j(v[1]);
        value.resize(v[1]);
        ok = read_exact(f, i.data(), v[1] * 8) && read_exact(f, j.data(), v[1] * 8) && read_exact(f, value.data(), v[1] * 4);
        flat_index.resize(v[1]);
        for (uint64_t k = 0; ok && k < v[1]; ++k) flat_index[k] = i[k] * nvoxel + j[k];  // source/laplacian.cpp:67
    }
    std::fclose(f);
    return ok;
}

}  // namespace rawfiles
