// The reference's HDF5 file schemas read through h5lite: what categorize_input_files,
// sort_rtm_files, the consistency checks, get_total_rtm_size, read_rtm_frame_masks (source/hdf5files.cpp),
// RayTransferMatrix::read_hdf5 (source/raytransfer.cpp), CompositeImage's readers (source/image.cpp:53-72, 268-314),
// LaplacianMatrix::read_hdf5 (source/laplacian.cpp) and the voxel-grid pass-through (source/voxelgrid.cpp:41-187) do.
#pragma once

#include <cstdint>
#include <map>
#include <string>
#include <vector>

#include "frame_source.hpp"
#include "h5lite.hpp"

namespace h5schema {

bool is_hdf5(const std::string& path);

struct RtmBlock {  // one camera x voxel-segment file
    std::string path, camera, dataset;  // dataset = "rtm/<name>/value"
    uint64_t pixel_offset = 0, npixel = 0, voxel_offset = 0, nvoxel = 0;
    bool sparse = false;
    double wavelength = 0;
};

struct VoxelGrid {
    uint64_t nx = 0, ny = 0, nz = 0;
    double xmin = 0, xmax = 1, ymin = 0, ymax = 1, zmin = 0, zmax = 1;
    std::string coordinate_system = "cartesian";
    std::vector<int32_t> voxmap;  // [nx * ny * nz], -1 = no voxel
};

struct Problem {
    std::vector<RtmBlock> blocks;          // camera-major (camera name order), segments by lowest voxel-map cell
    std::vector<CameraSeries> cameras;     // camera name order; pixels = the frame-mask pixels in row-major order
    std::vector<std::string> camera_names;
    uint64_t npixel = 0, nvoxel = 0;
    VoxelGrid grid;
};

// Throws std::runtime_error with the reference's wording where it has one.
Problem open_problem(const std::vector<std::string>& input_files, const std::string& rtm_name, double wavelength_threshold);

// rows [row0, row0 + nrows) of a block as dense fp32 [nrows][block.nvoxel] (sparse blocks are scattered into it)
std::vector<float> read_block_rows(const RtmBlock& block, uint64_t row0, uint64_t nrows);

void read_laplacian(const std::string& path, uint64_t nvoxel, std::vector<uint64_t>& flat_index, std::vector<float>& value);

// The reference's output file (source/solution.cpp:60-165, source/voxelgrid.cpp:112-187), written once.
// The reference's solution file (source/solution.cpp): created once, extended every few frames, complete on disk after
// every append. Host memory holds only the frames since the last append.
class SolutionFile {
public:
    SolutionFile(const std::string& path, const std::vector<std::string>& camera_names, uint64_t nvoxel,
                 uint64_t frames_per_flush, const VoxelGrid& grid);
    ~SolutionFile();
    SolutionFile(const SolutionFile&) = delete;
    SolutionFile& operator=(const SolutionFile&) = delete;
    void append(const std::vector<double>& values /* [n][nvoxel] */, const std::vector<double>& time,
                const std::vector<std::vector<double>>& camera_time /* [n][ncamera] */, const std::vector<int32_t>& status,
                const std::vector<int32_t>& iterations);

private:
    h5lite::Writer* writer;
    std::vector<std::string> cameras;
    uint64_t nvox;
};

}  // namespace h5schema
