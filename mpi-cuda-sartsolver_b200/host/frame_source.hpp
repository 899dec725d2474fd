// Frame source of the driver: which frames of which camera make up each composite frame, and reading them.
// A restatement of the reference's CompositeImage time logic (source/image.cpp:53-107 selection by time interval,
// :110-196 multi-camera synchronisation on a time grid) over the raw image files of raw_files.hpp. Caching of frames
// (source/image.cpp:268-331, max_cache_size frames read ahead) becomes read_chunk(): the frames of one solver call are
// fetched with ONE read per camera and run of consecutive frame indices, straight into the caller's chunk buffer.
#pragma once

#include <array>
#include <cstdint>
#include <string>
#include <vector>

#include <functional>

#include "raw_files.hpp"

// One camera's image series: where its pixels sit in the composite frame, its timeline, and how to read pixels
// [first, first + count) of its frame `index` (the container -- flat file or HDF5 -- is the caller's business).
struct CameraSeries {
    std::string name;
    uint64_t pixel_offset = 0, npixel = 0;
    std::vector<double> time;
    std::function<bool(size_t index, uint64_t first, uint64_t count, double* out, std::string& error)> read;
    // optional: pixels [first, first + count) of the frames index .. index + nframes - 1 in one go; frame f goes to
    // out + f * out_stride. Absent: read() frame by frame.
    std::function<bool(size_t index, size_t nframes, uint64_t first, uint64_t count, double* out, uint64_t out_stride,
                       std::string& error)> read_run;
};

CameraSeries camera_from_raw(const rawfiles::ImgHeader& header);

class CompositeFrames {
public:
    // cameras in composite order (the reference: lexicographic camera name); their pixel ranges must tile [0, npixel)
    CompositeFrames(std::vector<CameraSeries> cameras, const std::vector<std::array<double, 4>>& time_intervals);
    const std::vector<CameraSeries>& cameras() const { return images; }

    size_t size() const { return frame_indices.size(); }
    double frame_time(size_t k) const { return time[k]; }
    const std::vector<double>& camera_frame_time(size_t k) const { return camera_time[k]; }
    const std::vector<size_t>& indices(size_t k) const { return frame_indices[k]; }
    uint64_t npixel() const { return npix; }

    // pixels [offset, offset + count) of composite frame k
    bool read(size_t k, uint64_t offset, uint64_t count, std::vector<double>& out, std::string& error) const;
    // the same for composite frames k0 .. k0 + n - 1 at once: out = [n][count] (the reference's frame cache filled for a
    // chunk, source/image.cpp:268-331). *reads (optional) receives the number of read requests issued.
    bool read_chunk(size_t k0, size_t n, uint64_t offset, uint64_t count, std::vector<double>& out, std::string& error,
                    size_t* reads = nullptr) const;

    // the synchronisation step on its own (exposed for tests): per camera (time, frame index) pairs inside one interval
    void add_interval(const std::vector<std::vector<std::pair<double, size_t>>>& timepairs, double step, double threshold);

private:
    std::vector<CameraSeries> images;
    uint64_t npix = 0;
    std::vector<std::vector<size_t>> frame_indices;  // [composite frame][camera] -> frame index in that camera's file
    std::vector<std::vector<double>> camera_time;
    std::vector<double> time;
};
