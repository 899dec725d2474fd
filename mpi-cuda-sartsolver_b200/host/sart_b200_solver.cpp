// `sartsolver` for machines without libhdf5: the reference's main() (source/main.cpp:25-151) over the flat files of
// raw_files.hpp -- the same command line (arguments_b200.cpp), the same input categorisation (RTM vs image files, any
// order: source/hdf5files.cpp:20-43), matrix extents from the files (:349-389), frame selection and multi-camera
// synchronisation (frame_source.cpp), Laplacian, solver set-up, frame loop with warm start / --no_guess, per-frame
// "Processed in: ... ms", and one solution record per frame. The matrix never sits in host memory: every file block is
// streamed into the GPU shard through sart_load_rows_file's pinned staging ring. Single process / single GPU; a multi-GPU
// job is the same program per rank with the communicator id shipped as shown in INTEGRATION.md.
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <iostream>
#include <limits>
#include <stdexcept>
#include <vector>

#include "../../include/sart_b200.h"
#include "arguments_b200.hpp"
#include "frame_source.hpp"
#include "raw_files.hpp"

namespace {

[[noreturn]] void die(const std::string& message) {
    std::cerr << message << std::endl;
    std::exit(1);
}

void must(int rc) {
    if (rc != SART_OK) die(sart_last_error());
}

}  // namespace

int main(int argc, char* argv[]) {
    const Config config = parse_arguments_or_exit(argc, argv);
    std::vector<std::array<double, 4>> time_intervals;
    std::string error;
    if (!parse_time_intervals(config.time_range, time_intervals, error)) die(error);
    if (config.use_cpu) die("This build has no CPU solver (--use_cpu): the engine runs on the GPU only.");

    // source/hdf5files.cpp:20-43: every input is either a ray-transfer file or an image file
    std::vector<rawfiles::RtmHeader> matrix_files;
    std::vector<rawfiles::ImgHeader> image_files;
    for (const auto& path : config.input_files) {
        const std::string magic = rawfiles::file_magic(path);
        if (magic == rawfiles::kRtmMagic) {
            rawfiles::RtmHeader h;
            if (!rawfiles::read_rtm_header(path, h)) die("Unable to read " + path + ".");
            matrix_files.push_back(h);
        } else if (magic == rawfiles::kImgMagic) {
            rawfiles::ImgHeader h;
            if (!rawfiles::read_img_header(path, h)) die("Unable to read " + path + ".");
            image_files.push_back(h);
        } else {
            die("File " + path + " is neither a ray-transfer matrix nor an image file.");
        }
    }
    if (matrix_files.empty() || image_files.empty()) die("At least one ray-transfer file and one image file are required.");

    // source/hdf5files.cpp:349-389: total size from the files; the blocks must cover it exactly once
    uint64_t npixel = 0, nvoxel = 0, covered = 0;
    for (const auto& h : matrix_files) {
        npixel = std::max(npixel, h.pixel_offset + h.npixel);
        nvoxel = std::max(nvoxel, h.voxel_offset + h.nvoxel);
        covered += h.npixel * h.nvoxel;
    }
    if (covered != npixel * nvoxel) die("Ray-transfer files do not tile the matrix exactly once.");

    CompositeFrames* frames = nullptr;
    try {
        frames = new CompositeFrames(image_files, time_intervals);
    } catch (const std::runtime_error& err) {
        die(err.what());
    }
    if (frames->npixel() != npixel) die("Ray-transfer matrices and images have different number of pixels.");

    const int rank = 0, numproc = 1;
    sart_config cfg;
    must(sart_config_default(&cfg));
    cfg.npixel = npixel;
    cfg.nvoxel = nvoxel;
    cfg.rank = rank;
    cfg.nranks = numproc;
    cfg.logarithmic = config.logarithmic ? 1 : 0;
    cfg.ray_density_threshold = config.ray_density_threshold;  // source/main.cpp:105-110
    cfg.ray_length_threshold = config.ray_length_threshold;
    cfg.conv_tolerance = config.conv_tolerance;
    cfg.beta_laplace = config.beta_laplace;
    cfg.relaxation = config.relaxation;
    cfg.max_iterations = config.max_iterations;
    sart_engine* engine = nullptr;
    must(sart_create(&cfg, &engine));
    uint64_t offset_pixel = 0, npixel_local = 0;
    must(sart_shard_range(npixel, numproc, rank, &offset_pixel, &npixel_local));

    // source/raytransfer.cpp:27-127 without the host copy: each file block's rows that fall into this rank's range
    const auto t_load = std::chrono::steady_clock::now();
    for (const auto& h : matrix_files) {
        const uint64_t lo = std::max(h.pixel_offset, offset_pixel), hi = std::min(h.pixel_offset + h.npixel, offset_pixel + npixel_local);
        if (lo >= hi) continue;
        const uint64_t file_offset = rawfiles::RtmHeader::kDataOffset + (lo - h.pixel_offset) * h.nvoxel * sizeof(float);
        must(sart_load_rows_file(engine, h.path.c_str(), file_offset, h.nvoxel, lo - offset_pixel, hi - lo, h.voxel_offset, h.nvoxel));
    }
    const std::chrono::duration<double, std::milli> load_ms(std::chrono::steady_clock::now() - t_load);

    if (!config.laplacian_file.empty()) {  // source/main.cpp:73-74
        uint64_t lap_nvoxel = 0;
        std::vector<uint64_t> index;
        std::vector<float> value;
        if (!rawfiles::read_laplacian(config.laplacian_file, lap_nvoxel, index, value)) die("Unable to read " + config.laplacian_file + ".");
        if (lap_nvoxel != nvoxel) die("Laplacian and ray-transfer matrices have different number of voxels.");  // laplacian.cpp:46-49
        must(sart_set_laplacian_coo(engine, index.data(), value.data(), index.size()));
    }
    must(sart_finalize(engine));
    std::cout << "Matrix " << npixel << " x " << nvoxel << " loaded in: " << load_ms.count() << " ms" << std::endl;

    FILE* out = std::fopen(config.output_file.c_str(), "wb");
    if (!out) die("Unable to open " + config.output_file + ".");
    const uint64_t head[2] = {nvoxel, (uint64_t)frames->size()};
    std::fwrite(rawfiles::kSolMagic, 1, 8, out);
    std::fwrite(head, sizeof(uint64_t), 2, out);

    // source/main.cpp:127-140
    std::vector<double> solution_vec, frame;
    for (size_t k = 0; k < frames->size(); ++k) {
        if (!frames->read(k, offset_pixel, npixel_local, frame, error)) die(error);
        const auto last = std::chrono::steady_clock::now();
        const int has_guess = solution_vec.empty() ? 0 : 1;
        if (!has_guess) solution_vec.resize(nvoxel);
        int32_t status = 0, iterations = 0;
        must(sart_solve(engine, frame.data(), solution_vec.data(), has_guess, &status, &iterations));
        const std::chrono::duration<double, std::milli> duration(std::chrono::steady_clock::now() - last);
        std::cout << "Processed in: " << duration.count() << " ms" << std::endl;
        const double t = frames->frame_time(k);
        std::fwrite(&t, sizeof(double), 1, out);
        std::fwrite(&status, sizeof(int32_t), 1, out);
        std::fwrite(&iterations, sizeof(int32_t), 1, out);
        std::fwrite(solution_vec.data(), sizeof(double), nvoxel, out);
        if (config.no_guess) solution_vec.clear();
    }
    std::fclose(out);
    sart_destroy(engine);
    delete frames;
    return 0;
}
