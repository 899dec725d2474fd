// Frame source of the driver: which frames of which camera make up each composite frame, and reading them.
// A restatement of the reference's CompositeImage time logic (source/image.cpp:53-107 selection by time interval,
// :110-196 multi-camera synchronisation on a time grid) over the raw image files of raw_files.hpp. Caching of frames
// (source/image.cpp:268-331) is not needed here: a frame is read when it is solved.
#pragma once

#include <array>
#include <cstdint>
#include <string>
#include <vector>

#include "raw_files.hpp"

class CompositeFrames {
public:
    // images: one per camera, ordered by pixel_offset; their pixel ranges must tile [0, npixel)
    CompositeFrames(std::vector<rawfiles::ImgHeader> images, const std::vector<std::array<double, 4>>& time_intervals);

    size_t size() const { return frame_indices.size(); }
    double frame_time(size_t k) const { return time[k]; }
    const std::vector<double>& camera_frame_time(size_t k) const { return camera_time[k]; }
    const std::vector<size_t>& indices(size_t k) const { return frame_indices[k]; }
    uint64_t npixel() const { return npix; }

    // pixels [offset, offset + count) of composite frame k
    bool read(size_t k, uint64_t offset, uint64_t count, std::vector<double>& out, std::string& error) const;

    // the synchronisation step on its own (exposed for tests): per camera (time, frame index) pairs inside one interval
    void add_interval(const std::vector<std::vector<std::pair<double, size_t>>>& timepairs, double step, double threshold);

private:
    std::vector<rawfiles::ImgHeader> images;
    uint64_t npix = 0;
    std::vector<std::vector<size_t>> frame_indices;  // [composite frame][camera] -> frame index in that camera's file
    std::vector<std::vector<double>> camera_time;
    std::vector<double> time;
};
