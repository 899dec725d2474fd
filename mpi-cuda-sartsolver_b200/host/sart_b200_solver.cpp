// `sartsolver` for machines without libhdf5: the reference's main() (source/main.cpp:25-151) over the flat files of
// raw_files.hpp -- the same command line (arguments_b200.cpp), the same input categorisation (RTM vs image files, any
// order: source/hdf5files.cpp:20-43), matrix extents from the files (:349-389), frame selection and multi-camera
// synchronisation (frame_source.cpp), Laplacian, solver set-up, frame loop with warm start / --no_guess, per-frame
// "Processed in: ... ms", and one solution record per frame. The matrix never sits in host memory: every file block is
// streamed into the GPU shard through sart_load_rows_file's pinned staging ring.
// Several GPUs: start one process per GPU with RANK / WORLD_SIZE (torchrun style) or OMPI_COMM_WORLD_RANK / _SIZE or
// PMI_RANK / PMI_SIZE in the environment (what mpirun / srun export); since there is no MPI library to link against on
// this image, the 128-byte communicator id travels from rank 0 to the others through the file named by
// SART_COMM_ID_FILE (default /tmp/sart_b200_comm_id.<MASTER_PORT or parent pid>). Every rank loads only its own rows
// and pixels; rank 0 alone prints and writes the solution, as in the reference.
#include <unistd.h>

#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <thread>
#include <cstdio>
#include <iostream>
#include <limits>
#include <stdexcept>
#include <vector>

#include "../../include/sart_b200.h"
#include "arguments_b200.hpp"
#include "frame_source.hpp"
#include "h5_schema.hpp"
#include "raw_files.hpp"

namespace {

[[noreturn]] void die(const std::string& message) {
    std::cerr << message << std::endl;
    std::exit(1);
}

void must(int rc) {
    if (rc != SART_OK) die(sart_last_error());
}

int env_int(std::initializer_list<const char*> names, int dflt) {
    for (const char* n : names)
        if (const char* v = std::getenv(n))
            if (*v) return std::atoi(v);
    return dflt;
}

// rank 0 publishes the communicator id through a file (write + atomic rename), the others wait for it
void exchange_comm_id(int rank, uint8_t id[SART_UNIQUE_ID_BYTES], std::string& path) {
    if (const char* p = std::getenv("SART_COMM_ID_FILE")) path = p;
    else {
        const char* port = std::getenv("MASTER_PORT");
        path = "/tmp/sart_b200_comm_id." + (port ? std::string(port) : std::to_string((long)getppid()));
    }
    if (rank == 0) {
        must(sart_comm_unique_id(id));
        const std::string tmp = path + ".tmp";
        FILE* f = std::fopen(tmp.c_str(), "wb");
        if (!f || std::fwrite(id, 1, SART_UNIQUE_ID_BYTES, f) != SART_UNIQUE_ID_BYTES) die("Unable to write " + tmp + ".");
        std::fclose(f);
        if (std::rename(tmp.c_str(), path.c_str()) != 0) die("Unable to publish " + path + ".");
        return;
    }
    for (int waited_ms = 0; waited_ms < 120000; waited_ms += 50) {
        FILE* f = std::fopen(path.c_str(), "rb");
        if (f) {
            const bool ok = std::fread(id, 1, SART_UNIQUE_ID_BYTES, f) == SART_UNIQUE_ID_BYTES;
            std::fclose(f);
            if (ok) return;
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(50));
    }
    die("Timed out waiting for the communicator id in " + path + ".");
}

}  // namespace

int main(int argc, char* argv[]) {
    const Config config = parse_arguments_or_exit(argc, argv);
    std::vector<std::array<double, 4>> time_intervals;
    std::string error;
    if (!parse_time_intervals(config.time_range, time_intervals, error)) die(error);

    // Inputs are either all HDF5 files in the reference's schemas (read with h5lite) or all flat stand-in files.
    const bool hdf5 = h5schema::is_hdf5(config.input_files.front());
    for (const auto& path : config.input_files)
        if (h5schema::is_hdf5(path) != hdf5) die("HDF5 and flat input files cannot be mixed.");

    std::vector<rawfiles::RtmHeader> matrix_files;
    h5schema::Problem h5;
    std::vector<CameraSeries> cameras;
    std::vector<std::string> camera_names;
    uint64_t npixel = 0, nvoxel = 0;
    if (hdf5) {
        try {
            h5 = h5schema::open_problem(config.input_files, config.raytransfer_name, config.wavelength_threshold);
        } catch (const std::exception& err) {
            die(err.what());
        }
        npixel = h5.npixel;
        nvoxel = h5.nvoxel;
        cameras = h5.cameras;
        camera_names = h5.camera_names;
    } else {
        // source/hdf5files.cpp:20-43: every input is either a ray-transfer file or an image file
        for (const auto& path : config.input_files) {
            const std::string magic = rawfiles::file_magic(path);
            if (magic == rawfiles::kRtmMagic) {
                rawfiles::RtmHeader h;
                if (!rawfiles::read_rtm_header(path, h)) die("Unable to read " + path + ".");
                matrix_files.push_back(h);
            } else if (magic == rawfiles::kImgMagic) {
                rawfiles::ImgHeader h;
                if (!rawfiles::read_img_header(path, h)) die("Unable to read " + path + ".");
                cameras.push_back(camera_from_raw(h));
                camera_names.push_back("camera" + std::to_string(camera_names.size()));
            } else {
                die("File " + path + " is neither a ray-transfer matrix nor an image file.");
            }
        }
        if (matrix_files.empty() || cameras.empty()) die("At least one ray-transfer file and one image file are required.");
        // source/hdf5files.cpp:349-389: total size from the files; the blocks must cover it exactly once
        uint64_t covered = 0;
        for (const auto& h : matrix_files) {
            npixel = std::max(npixel, h.pixel_offset + h.npixel);
            nvoxel = std::max(nvoxel, h.voxel_offset + h.nvoxel);
            covered += h.npixel * h.nvoxel;
        }
        if (covered != npixel * nvoxel) die("Ray-transfer files do not tile the matrix exactly once.");
    }

    CompositeFrames* frames = nullptr;
    try {
        frames = new CompositeFrames(cameras, time_intervals);
    } catch (const std::runtime_error& err) {
        die(err.what());
    }
    if (frames->npixel() != npixel) die("Ray-transfer matrices and images have different number of pixels.");

    // source/main.cpp:63-65 (MPI_Comm_rank / MPI_Comm_size)
    const int rank = env_int({"RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK"}, 0);
    const int numproc = env_int({"WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE"}, 1);
    if (numproc < 1 || rank < 0 || rank >= numproc) die("Inconsistent rank / world size in the environment.");
    sart_config cfg;
    must(sart_config_default(&cfg));
    cfg.npixel = npixel;
    cfg.nvoxel = nvoxel;
    cfg.rank = rank;
    cfg.nranks = numproc;
    cfg.logarithmic = config.logarithmic ? 1 : 0;
    // Which of the reference's two solvers this command line would have run decides the arithmetic where they differ
    // observably (saturated pixels in the initial guess, the 1e-7 floor on the starting solution, the log solver's epsilon
    // and penalty units): without --use_cpu the reference runs its CUDA solver (source/main.cpp:88-97),
    // with it the CPU solver. Everything still runs on the GPU -- --use_cpu selects the CPU solver's semantics, not a CPU.
    cfg.semantics = config.use_cpu ? SART_SEMANTICS_CPU : SART_SEMANTICS_GPU;
    cfg.ray_density_threshold = config.ray_density_threshold;  // source/main.cpp:105-110
    cfg.ray_length_threshold = config.ray_length_threshold;
    cfg.conv_tolerance = config.conv_tolerance;
    cfg.beta_laplace = config.beta_laplace;
    cfg.relaxation = config.relaxation;
    cfg.max_iterations = config.max_iterations;
    cfg.device = env_int({"LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK"}, -1);
    sart_engine* engine = nullptr;
    must(sart_create(&cfg, &engine));
    std::string id_file;
    if (numproc > 1) {
        uint8_t id[SART_UNIQUE_ID_BYTES];
        exchange_comm_id(rank, id, id_file);
        must(sart_comm_init(engine, id));
        if (rank == 0) std::remove(id_file.c_str());  // every rank has joined once sart_comm_init returns
    }
    uint64_t offset_pixel = 0, npixel_local = 0;
    must(sart_shard_range(npixel, numproc, rank, &offset_pixel, &npixel_local));

    // source/raytransfer.cpp:27-127 without the host copy: each file block's rows that fall into this rank's range
    const auto t_load = std::chrono::steady_clock::now();
    for (const auto& b : h5.blocks) {
        const uint64_t lo = std::max(b.pixel_offset, offset_pixel), hi = std::min(b.pixel_offset + b.npixel, offset_pixel + npixel_local);
        if (lo >= hi) continue;
        try {
            h5lite::File f(b.path);
            const uint64_t data_offset = b.sparse ? UINT64_MAX : f.contiguous_offset(b.dataset);
            if (data_offset != UINT64_MAX) {
                // contiguous, unfiltered dataset: stream it straight from the file (what H5Dget_offset allows)
                must(sart_load_rows_file(engine, b.path.c_str(), data_offset + (lo - b.pixel_offset) * b.nvoxel * sizeof(float), b.nvoxel,
                                         lo - offset_pixel, hi - lo, b.voxel_offset, b.nvoxel));
            } else {
                // chunked / compressed / sparse: decode slabs of rows on the host
                const uint64_t slab = std::max<uint64_t>(1, (64ull << 20) / (b.nvoxel * sizeof(float)));
                for (uint64_t r = lo; r < hi; r += slab) {
                    const uint64_t n = std::min(slab, hi - r);
                    const std::vector<float> rows = h5schema::read_block_rows(b, r - b.pixel_offset, n);
                    must(sart_upload_rows(engine, r - offset_pixel, n, b.voxel_offset, b.nvoxel, rows.data(), b.nvoxel));
                }
            }
        } catch (const std::exception& err) {
            die(err.what());
        }
    }
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
        if (h5schema::is_hdf5(config.laplacian_file)) {
            try {
                h5schema::read_laplacian(config.laplacian_file, nvoxel, index, value);
                lap_nvoxel = nvoxel;
            } catch (const std::exception& err) {
                die(err.what());
            }
        } else if (!rawfiles::read_laplacian(config.laplacian_file, lap_nvoxel, index, value)) {
            die("Unable to read " + config.laplacian_file + ".");
        }
        if (lap_nvoxel != nvoxel) die("Laplacian and ray-transfer matrices have different number of voxels.");  // laplacian.cpp:46-49
        must(sart_set_laplacian_coo(engine, index.data(), value.data(), index.size()));
    }
    must(sart_finalize(engine));
    if (rank == 0) std::cout << "Matrix " << npixel << " x " << nvoxel << " loaded in: " << load_ms.count() << " ms" << std::endl;

    // flat output: one record per frame as it is solved; HDF5 output: the reference's solution file, written at the end
    const bool h5_out = config.output_file.size() > 3 && (config.output_file.rfind(".h5") == config.output_file.size() - 3 ||
                                                          config.output_file.rfind(".hdf5") == config.output_file.size() - 5);
    FILE* out = nullptr;
    if (!h5_out && rank == 0) {
        out = std::fopen(config.output_file.c_str(), "wb");
        if (!out) die("Unable to open " + config.output_file + ".");
        const uint64_t head[2] = {nvoxel, (uint64_t)frames->size()};
        std::fwrite(rawfiles::kSolMagic, 1, 8, out);
        std::fwrite(head, sizeof(uint64_t), 2, out);
    }
    std::vector<double> all_values, all_time;
    std::vector<std::vector<double>> all_camera_time;
    std::vector<int32_t> all_status, all_iterations;

    // source/main.cpp:127-140. With --no_guess consecutive frames do not depend on each other, so they are handed to the
    // engine eight at a time (sart_solve_batch: every pass over the matrix serves the whole group).
    auto store = [&](size_t k, const double* x, int32_t status, int32_t iterations) {
        const double t = frames->frame_time(k);
        if (rank != 0) return;  // source/main.cpp:134: only the root stores solutions
        if (h5_out) {
            all_values.insert(all_values.end(), x, x + nvoxel);
            all_time.push_back(t);
            all_camera_time.push_back(frames->camera_frame_time(k));
            all_status.push_back(status);
            all_iterations.push_back(iterations);
        } else {
            std::fwrite(&t, sizeof(double), 1, out);
            std::fwrite(&status, sizeof(int32_t), 1, out);
            std::fwrite(&iterations, sizeof(int32_t), 1, out);
            std::fwrite(x, sizeof(double), nvoxel, out);
        }
    };
    h5schema::SolutionFile* h5_file = nullptr;
    auto flush = [&](size_t done) {
        // source/solution.cpp:55, 114-180: the cached solutions are appended to the file every max_cached_solutions frames
        // (and at the end), so a long run keeps at most that many in host memory and leaves a complete file behind
        // whenever it stops
        if (!(h5_out && rank == 0) || all_time.empty()) return;
        if (done < frames->size() && all_time.size() < (size_t)config.max_cached_solutions) return;
        try {
            if (!h5_file)
                h5_file = new h5schema::SolutionFile(config.output_file, camera_names, nvoxel, (uint64_t)config.max_cached_solutions, h5.grid);
            h5_file->append(all_values, all_time, all_camera_time, all_status, all_iterations);
        } catch (const std::exception& err) {
            die(err.what());
        }
        all_values.clear();
        all_time.clear();
        all_camera_time.clear();
        all_status.clear();
        all_iterations.clear();
    };
    std::vector<double> solution_vec, frame;
    size_t k = 0;
    while (k < frames->size()) {
        const size_t nb = config.no_guess ? std::min<size_t>(8, frames->size() - k) : 1;
        if (nb >= 2) {
            std::vector<double> group_frames, group_solutions(nb * nvoxel);
            if (!frames->read_chunk(k, nb, offset_pixel, npixel_local, group_frames, error)) die(error);
            const auto last = std::chrono::steady_clock::now();
            std::vector<int32_t> status(nb, 0), iterations(nb, 0);
            must(sart_solve_batch(engine, (int32_t)nb, group_frames.data(), group_solutions.data(), status.data(), iterations.data()));
            const std::chrono::duration<double, std::milli> duration(std::chrono::steady_clock::now() - last);
            for (size_t b = 0; b < nb; ++b) {
                if (rank == 0) std::cout << "Processed in: " << duration.count() / (double)nb << " ms" << std::endl;
                store(k + b, group_solutions.data() + b * nvoxel, status[b], iterations[b]);
                flush(k + b + 1);
            }
        } else if (!config.no_guess && frames->size() - k >= 2) {
            // warm-started frames in chunks of max_cached_frames (the reference's frame cache, source/image.cpp:268-331):
            // sart_solve_series knows the next frame while it solves the current one
            const size_t nc = std::min<size_t>((size_t)config.max_cached_frames, frames->size() - k);
            // (one read per camera and run of consecutive frames, like the reference's cache refill)
            std::vector<double> chunk, chunk_solutions(nc * nvoxel);
            if (!frames->read_chunk(k, nc, offset_pixel, npixel_local, chunk, error)) die(error);
            const auto last = std::chrono::steady_clock::now();
            std::vector<int32_t> status(nc, 0), iterations(nc, 0);
            must(sart_solve_series(engine, (int32_t)nc, chunk.data(), solution_vec.empty() ? nullptr : solution_vec.data(),
                                   chunk_solutions.data(), status.data(), iterations.data()));
            const std::chrono::duration<double, std::milli> duration(std::chrono::steady_clock::now() - last);
            for (size_t b = 0; b < nc; ++b) {
                if (rank == 0) std::cout << "Processed in: " << duration.count() / (double)nc << " ms" << std::endl;
                store(k + b, chunk_solutions.data() + b * nvoxel, status[b], iterations[b]);
                flush(k + b + 1);
            }
            solution_vec.assign(chunk_solutions.end() - (std::ptrdiff_t)nvoxel, chunk_solutions.end());
            k += nc;
            continue;
        } else {
            if (!frames->read(k, offset_pixel, npixel_local, frame, error)) die(error);
            const auto last = std::chrono::steady_clock::now();
            const int has_guess = solution_vec.empty() ? 0 : 1;
            if (!has_guess) solution_vec.resize(nvoxel);
            int32_t status = 0, iterations = 0;
            must(sart_solve(engine, frame.data(), solution_vec.data(), has_guess, &status, &iterations));
            const std::chrono::duration<double, std::milli> duration(std::chrono::steady_clock::now() - last);
            if (rank == 0) std::cout << "Processed in: " << duration.count() << " ms" << std::endl;
            store(k, solution_vec.data(), status, iterations);
            if (config.no_guess) solution_vec.clear();
            flush(k + 1);
        }
        k += nb;
    }
    if (h5_out && rank == 0) {
        flush(frames->size());
        delete h5_file;
    } else if (out) {
        std::fclose(out);
    }
    sart_destroy(engine);
    delete frames;
    return 0;
}
