#include "h5_schema.hpp"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <memory>
#include <stdexcept>

namespace h5schema {

bool is_hdf5(const std::string& path) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    const uint8_t sig[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
    uint8_t buf[8];
    bool yes = false;
    for (long off = 0; off < (1L << 20) && !yes; off = off ? off * 2 : 512) {
        if (std::fseek(f, off, SEEK_SET) != 0 || std::fread(buf, 1, 8, f) != 8) break;
        yes = std::memcmp(buf, sig, 8) == 0;
    }
    std::fclose(f);
    return yes;
}

namespace {

struct RtmFile {
    std::string path, camera;
    uint64_t npixel = 0, nvoxel = 0, min_cell = 0;
    double wavelength = 0;
    bool sparse = false;
    std::vector<uint8_t> mask;
    uint64_t mask_h = 0, mask_w = 0;
};

std::vector<uint8_t> read_mask(const h5lite::File& f, uint64_t& h, uint64_t& w) {
    const h5lite::Object o = f.object("rtm/frame_mask");
    if (o.dataset.dims.size() != 2) throw std::runtime_error("rtm/frame_mask of " + f.path() + " is not two-dimensional.");
    h = o.dataset.dims[0];
    w = o.dataset.dims[1];
    const auto v = f.read<int32_t>("rtm/frame_mask");
    std::vector<uint8_t> m(v.size());
    for (size_t i = 0; i < v.size(); ++i) m[i] = v[i] != 0;
    return m;
}

}  // namespace

Problem open_problem(const std::vector<std::string>& input_files, const std::string& rtm_name, double wavelength_threshold) {
    Problem pb;
    // source/hdf5files.cpp:20-43
    std::vector<std::string> matrix_files, image_files;
    for (const auto& name : input_files) {
        h5lite::File f(name);
        if (f.exists("rtm")) matrix_files.push_back(name);
        else if (f.exists("image")) image_files.push_back(name);
        else throw std::runtime_error("The file " + name + " is neither an RTM file nor an image file.");
    }
    if (matrix_files.empty() || image_files.empty())
        throw std::runtime_error("At least one RTM file and one image file are required.");

    // source/hdf5files.cpp:46-103: cameras by name, a camera's segment files by their lowest voxel-map cell
    std::map<std::string, std::map<uint64_t, RtmFile>> sorted;
    bool have_grid = false;
    double ref_wavelength = 0;
    for (const auto& name : matrix_files) {
        h5lite::File f(name);
        RtmFile r;
        r.path = name;
        r.camera = f.string_attribute("rtm", "camera_name");
        r.npixel = f.attribute<uint64_t>("rtm", "npixel");
        r.nvoxel = f.attribute<uint64_t>("rtm", "nvoxel");
        const std::string group = "rtm/" + rtm_name;
        if (!f.exists(group)) throw std::runtime_error("No group '" + group + "' in " + name + ".");
        r.wavelength = f.attribute<double>(group, "wavelength");
        r.sparse = f.attribute<int32_t>(group, "is_sparse") != 0;
        const uint64_t nx = f.attribute<uint64_t>("rtm/voxel_map", "nx"), ny = f.attribute<uint64_t>("rtm/voxel_map", "ny"),
                       nz = f.attribute<uint64_t>("rtm/voxel_map", "nz");
        if (!have_grid) {
            pb.grid.nx = nx;
            pb.grid.ny = ny;
            pb.grid.nz = nz;
            ref_wavelength = r.wavelength;
            have_grid = true;
        } else {
            // check_group_attribute_consistency, source/main.cpp:36-39
            if (nx != pb.grid.nx || ny != pb.grid.ny || nz != pb.grid.nz)
                throw std::runtime_error("Attribute nx, ny or nz of rtm/voxel_map differs between the RTM files.");
            if (r.wavelength != ref_wavelength)
                throw std::runtime_error("Attribute wavelength of rtm/" + rtm_name + " differs between the RTM files.");
        }
        const auto i = f.read<uint64_t>("rtm/voxel_map/i"), j = f.read<uint64_t>("rtm/voxel_map/j"), k = f.read<uint64_t>("rtm/voxel_map/k");
        r.min_cell = nx * ny * nz;
        for (size_t q = 0; q < i.size(); ++q) r.min_cell = std::min(r.min_cell, i[q] * ny * nz + j[q] * nz + k[q]);
        r.mask = read_mask(f, r.mask_h, r.mask_w);
        sorted[r.camera][r.min_cell] = r;
    }

    // per camera: frame masks equal over segments (:106-143), pixel count = mask pixels, blocks with global offsets
    std::map<std::string, const RtmFile*> first_of_camera;
    std::vector<int32_t> ref_voxmap;
    uint64_t pixel_offset = 0;
    for (const auto& cam : sorted) {
        const RtmFile& first = cam.second.begin()->second;
        first_of_camera[cam.first] = &first;
        uint64_t voxel_offset = 0;
        std::vector<int32_t> voxmap(pb.grid.nx * pb.grid.ny * pb.grid.nz, -1);
        int64_t nsource_prev = 0;
        for (const auto& seg : cam.second) {
            const RtmFile& r = seg.second;
            if (r.mask != first.mask) throw std::runtime_error("RTM files for " + cam.first + " view have different frame masks.");
            if (r.npixel != first.npixel) throw std::runtime_error("RTM files for " + cam.first + " view have different number of pixels.");
            RtmBlock b;
            b.path = r.path;
            b.camera = cam.first;
            b.dataset = "rtm/" + rtm_name + "/value";
            b.pixel_offset = pixel_offset;
            b.npixel = r.npixel;
            b.voxel_offset = voxel_offset;
            b.nvoxel = r.nvoxel;
            b.sparse = r.sparse;
            b.wavelength = r.wavelength;
            pb.blocks.push_back(b);
            voxel_offset += r.nvoxel;
            // voxel map of this camera (check_rtm_voxel_consistency :146-218, BaseVoxelGrid::read_hdf5)
            h5lite::File f(r.path);
            const auto i = f.read<uint64_t>("rtm/voxel_map/i"), j = f.read<uint64_t>("rtm/voxel_map/j"), k = f.read<uint64_t>("rtm/voxel_map/k");
            const auto value = f.read<int32_t>("rtm/voxel_map/value");
            for (size_t q = 0; q < i.size(); ++q) {
                const uint64_t cell = i[q] * pb.grid.ny * pb.grid.nz + j[q] * pb.grid.nz + k[q];
                if (cell >= voxmap.size()) throw std::runtime_error("Voxel map index out of range in " + r.path + ".");
                if (voxmap[cell] >= 0)
                    throw std::runtime_error("RTM segments for " + cam.first + " view have overlapping voxel maps at element (" +
                                             std::to_string(i[q]) + "," + std::to_string(j[q]) + "," + std::to_string(k[q]) + ").");
                voxmap[cell] = value[q] + (int32_t)nsource_prev;
            }
            nsource_prev += (int64_t)r.nvoxel;
        }
        if (ref_voxmap.empty()) {
            ref_voxmap = voxmap;
            pb.nvoxel = voxel_offset;
            h5lite::File f(first.path);
            auto opt = [&](const char* name, double dflt) { return f.has_attribute("rtm/voxel_map", name) ? f.attribute<double>("rtm/voxel_map", name) : dflt; };
            pb.grid.xmin = opt("xmin", 0);
            pb.grid.xmax = opt("xmax", 1);
            pb.grid.ymin = opt("ymin", 0);
            pb.grid.ymax = opt("ymax", 1);
            pb.grid.zmin = opt("zmin", 0);
            pb.grid.zmax = opt("zmax", 1);
            if (f.has_attribute("rtm/voxel_map", "coordinate_system")) {
                std::string cs = f.string_attribute("rtm/voxel_map", "coordinate_system");
                std::transform(cs.begin(), cs.end(), cs.begin(), [](unsigned char c) { return (char)std::tolower(c); });
                pb.grid.coordinate_system = (cs == "cylindrical") ? "cylindrical" : "cartesian";
            }
        } else {
            if (voxmap != ref_voxmap)
                throw std::runtime_error("RTM files for " + cam.first + " and " + sorted.begin()->first + " views have different voxel maps.");
            if (voxel_offset != pb.nvoxel) throw std::runtime_error("RTM files of different views have different number of voxels.");
        }
        uint64_t mask_pixels = 0;
        for (uint8_t m : first.mask) mask_pixels += m;
        if (mask_pixels != first.npixel)
            throw std::runtime_error("Frame mask of " + cam.first + " selects " + std::to_string(mask_pixels) + " pixels, the RTM has " +
                                     std::to_string(first.npixel) + ".");
        pixel_offset += first.npixel;
    }
    pb.npixel = pixel_offset;
    pb.grid.voxmap = ref_voxmap;

    // image files: one per camera (source/hdf5files.cpp sort_image_files, check_rtm_image_consistency :279-346)
    std::map<std::string, std::string> image_of_camera;
    for (const auto& name : image_files) {
        h5lite::File f(name);
        image_of_camera[f.string_attribute("image", "camera_name")] = name;
    }
    uint64_t offset = 0;
    for (const auto& cam : sorted) {
        auto it = image_of_camera.find(cam.first);
        if (it == image_of_camera.end()) throw std::runtime_error("No image file for " + cam.first + " view.");
        const RtmFile& r = *first_of_camera[cam.first];
        auto file = std::make_shared<h5lite::File>(it->second);
        const double wl = file->attribute<double>("image", "wavelength");
        if (std::abs(wl - r.wavelength) > wavelength_threshold)
            throw std::runtime_error("RTM and image wavelengths for " + cam.first + " view differ by more than the threshold.");
        const h5lite::Object fr = file->object("image/frame");
        if (fr.dataset.dims.size() != 3 || fr.dataset.dims[1] != r.mask_h || fr.dataset.dims[2] != r.mask_w)
            throw std::runtime_error("Image frames of " + it->second + " do not have the shape of the RTM frame mask.");
        CameraSeries c;
        c.name = cam.first;
        c.pixel_offset = offset;
        c.npixel = r.npixel;
        c.time = file->read<double>("image/time");
        if (c.time.size() != fr.dataset.dims[0]) throw std::runtime_error("image/time and image/frame of " + it->second + " differ in length.");
        // masked pixels in row-major order (source/image.cpp:278-314)
        auto selected = std::make_shared<std::vector<uint64_t>>();
        for (uint64_t p = 0; p < r.mask.size(); ++p)
            if (r.mask[p]) selected->push_back(p);
        c.read = [file, selected](size_t index, uint64_t first, uint64_t count, double* out, std::string& error) {
            try {
                const std::vector<double> frame = file->read_rows<double>("image/frame", index, 1);
                for (uint64_t q = 0; q < count; ++q) out[q] = frame[(*selected)[first + q]];
            } catch (const std::exception& err) {
                error = err.what();
                return false;
            }
            return true;
        };
        // a run of consecutive frames: one read of the dataset's rows [index, index + nframes), then the mask per frame
        c.read_run = [file, selected](size_t index, size_t nframes, uint64_t first, uint64_t count, double* out, uint64_t out_stride,
                                      std::string& error) {
            try {
                const std::vector<double> block = file->read_rows<double>("image/frame", index, nframes);
                const uint64_t frame_elems = block.size() / nframes;
                for (size_t k = 0; k < nframes; ++k)
                    for (uint64_t q = 0; q < count; ++q) out[k * out_stride + q] = block[k * frame_elems + (*selected)[first + q]];
            } catch (const std::exception& err) {
                error = err.what();
                return false;
            }
            return true;
        };
        pb.cameras.push_back(std::move(c));
        pb.camera_names.push_back(cam.first);
        offset += r.npixel;
    }
    return pb;
}

std::vector<float> read_block_rows(const RtmBlock& b, uint64_t row0, uint64_t nrows) {
    h5lite::File f(b.path);
    if (!b.sparse) return f.read_rows<float>(b.dataset, row0, nrows);
    // sparse triplets (source/raytransfer.cpp:67-91) scattered into the dense rows
    const std::string group = b.dataset.substr(0, b.dataset.rfind('/'));
    const auto pi = f.read<uint64_t>(group + "/pixel_index"), vi = f.read<uint64_t>(group + "/voxel_index");
    const auto val = f.read<float>(group + "/value");
    std::vector<float> dense(nrows * b.nvoxel, 0.f);
    for (size_t q = 0; q < val.size(); ++q)
        if (pi[q] >= row0 && pi[q] < row0 + nrows && vi[q] < b.nvoxel) dense[(pi[q] - row0) * b.nvoxel + vi[q]] = val[q];
    return dense;
}

void read_laplacian(const std::string& path, uint64_t nvoxel, std::vector<uint64_t>& flat_index, std::vector<float>& value) {
    h5lite::File f(path);
    if (f.attribute<uint64_t>("laplacian", "nvoxel") != nvoxel)
        throw std::runtime_error("Laplacian and ray-transfer matrices have different number of voxels.");  // laplacian.cpp:46-49
    const auto i = f.read<uint64_t>("laplacian/i"), j = f.read<uint64_t>("laplacian/j");
    value = f.read<float>("laplacian/value");
    if (i.size() != value.size() || j.size() != value.size()) throw std::runtime_error("Laplacian datasets differ in length.");
    flat_index.resize(value.size());
    for (size_t q = 0; q < value.size(); ++q) flat_index[q] = i[q] * nvoxel + j[q];
}

// Solution::create_hdf5 (source/solution.cpp:60-112): every dataset of the "solution" group has an unlimited first
// dimension; "value" is stored one frame per chunk, the one-dimensional ones in chunks of the first flush's length (here:
// of max_cached_solutions frames). The voxel map goes in once, when the file is created (source/main.cpp:143-145 writes it
// at the end; a file that is complete from the first flush on needs it then).
SolutionFile::SolutionFile(const std::string& path, const std::vector<std::string>& camera_names, uint64_t nvoxel,
                           uint64_t frames_per_flush, const VoxelGrid& g)
    : writer(new h5lite::Writer(path)), cameras(camera_names), nvox(nvoxel) {
    h5lite::Writer& w = *writer;
    const uint64_t chunk = std::max<uint64_t>(frames_per_flush, 1);
    w.create_group("solution");
    w.create_extendible("solution/value", true, {nvoxel}, 1);
    w.create_extendible("solution/time", true, {}, chunk);
    w.create_extendible("solution/status", false, {}, chunk);
    w.create_extendible("solution/iterations", false, {}, chunk);  // not in the reference's file
    for (const auto& name : cameras) w.create_extendible("solution/time_" + name, true, {}, chunk);
    if (!g.voxmap.empty()) {  // source/voxelgrid.cpp:112-187
        w.create_group("voxel_map");
        w.set_attribute("voxel_map", "nx", g.nx);
        w.set_attribute("voxel_map", "ny", g.ny);
        w.set_attribute("voxel_map", "nz", g.nz);
        w.set_attribute("voxel_map", "xmin", g.xmin);
        w.set_attribute("voxel_map", "xmax", g.xmax);
        w.set_attribute("voxel_map", "ymin", g.ymin);
        w.set_attribute("voxel_map", "ymax", g.ymax);
        w.set_attribute("voxel_map", "zmin", g.zmin);
        w.set_attribute("voxel_map", "zmax", g.zmax);
        w.set_attribute("voxel_map", "coordinate_system", g.coordinate_system);
        std::vector<int32_t> i, j, k, value;
        for (uint64_t cell = 0; cell < g.voxmap.size(); ++cell)
            if (g.voxmap[cell] > -1) {
                i.push_back((int32_t)(cell / (g.ny * g.nz)));
                const uint64_t rem = cell % (g.ny * g.nz);
                j.push_back((int32_t)(rem / g.nz));
                k.push_back((int32_t)(rem % g.nz));
                value.push_back(g.voxmap[cell]);
            }
        const uint64_t n = value.size();
        w.write_dataset("voxel_map/i", {n}, i.data());
        w.write_dataset("voxel_map/j", {n}, j.data());
        w.write_dataset("voxel_map/k", {n}, k.data());
        w.write_dataset("voxel_map/value", {n}, value.data());
    }
    w.close();  // the file exists from now on, with zero frames
}

SolutionFile::~SolutionFile() { delete writer; }

// Solution::update_hdf5 (source/solution.cpp:114-165): extend every dataset by the cached frames. After the call the file
// on disk is complete up to and including these frames.
void SolutionFile::append(const std::vector<double>& values, const std::vector<double>& time,
                          const std::vector<std::vector<double>>& camera_time, const std::vector<int32_t>& status,
                          const std::vector<int32_t>& iterations) {
    const uint64_t n = time.size();
    if (n == 0) return;
    if (values.size() != n * nvox || status.size() != n || iterations.size() != n || camera_time.size() != n)
        throw h5lite::Error("inconsistent solution cache");
    h5lite::Writer& w = *writer;
    w.append_rows("solution/value", values.data(), n);
    w.append_rows("solution/time", time.data(), n);
    w.append_rows("solution/status", status.data(), n);
    w.append_rows("solution/iterations", iterations.data(), n);
    for (size_t c = 0; c < cameras.size(); ++c) {
        std::vector<double> t(n);
        for (uint64_t f = 0; f < n; ++f) t[f] = camera_time[f][c];
        w.append_rows("solution/time_" + cameras[c], t.data(), n);
    }
    w.flush();
}

}  // namespace h5schema
