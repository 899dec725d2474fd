#include "frame_source.hpp"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <stdexcept>

namespace {
constexpr double kTimeEpsilon = 1.e-10;  // TIME_EPSILON of the reference (source/include/image.hpp)
}

CameraSeries camera_from_raw(const rawfiles::ImgHeader& h) {
    CameraSeries c;
    c.name = h.path;
    c.pixel_offset = h.pixel_offset;
    c.npixel = h.npixel;
    c.time = h.time;
    const std::string path = h.path;
    const uint64_t data_offset = h.data_offset(), npixel = h.npixel;
    c.read = [path, data_offset, npixel](size_t index, uint64_t first, uint64_t count, double* out, std::string& error) {
        FILE* f = std::fopen(path.c_str(), "rb");
        if (!f) {
            error = "cannot open " + path;
            return false;
        }
        const uint64_t pos = data_offset + (index * npixel + first) * 8;
        const bool ok = fseeko(f, (off_t)pos, SEEK_SET) == 0 && rawfiles::read_exact(f, out, count * 8);
        std::fclose(f);
        if (!ok) error = "short read on " + path;
        return ok;
    };
    return c;
}

CompositeFrames::CompositeFrames(std::vector<CameraSeries> cams, const std::vector<std::array<double, 4>>& time_intervals)
    : images(std::move(cams)) {
    std::sort(images.begin(), images.end(), [](const auto& a, const auto& b) { return a.pixel_offset < b.pixel_offset; });
    for (const auto& im : images) {
        if (im.pixel_offset != npix) throw std::runtime_error("Image files do not tile the pixel range: " + im.name);
        if (!std::is_sorted(im.time.begin(), im.time.end()))
            throw std::runtime_error("Image frames are not sorted by time in " + im.name + ".");
        npix += im.npixel;
    }
    // source/image.cpp:83-98: per interval, the frames of every camera with start <= t <= stop
    for (const auto& ti : time_intervals) {
        std::vector<std::vector<std::pair<double, size_t>>> timepairs;
        for (const auto& im : images) {
            std::vector<std::pair<double, size_t>> pairs;
            for (size_t i = 0; i < im.time.size(); ++i)
                if (im.time[i] >= ti[0] && im.time[i] <= ti[1]) pairs.emplace_back(im.time[i], i);
            timepairs.push_back(std::move(pairs));
        }
        add_interval(timepairs, ti[2], ti[3]);
    }
    if (frame_indices.empty()) throw std::runtime_error("No composite images can be created for given time intervals.");
}

// source/image.cpp:110-196. Frames of different cameras are combined when they fall on the same point of a time grid of
// pitch `step` within `threshold`; step 0 = the largest of the cameras' smallest frame spacings; threshold 0 = step.
void CompositeFrames::add_interval(const std::vector<std::vector<std::pair<double, size_t>>>& timepairs, double step,
                                   double threshold) {
    for (const auto& tp : timepairs)
        if (tp.empty()) return;  // a camera without frames in the interval: nothing can be combined
    double min_time = timepairs.front().front().first, max_time = timepairs.front().back().first;
    for (const auto& tp : timepairs) {
        min_time = std::min(tp.front().first, min_time);
        max_time = std::max(tp.back().first, max_time);
    }
    if (step == 0) {
        if (max_time - min_time < kTimeEpsilon) step = 1.;
        else
            for (const auto& tp : timepairs) {
                double min_diff = tp.back().first - tp.front().first;
                for (size_t i = 0; i + 1 < tp.size(); ++i) min_diff = std::min(tp[i + 1].first - tp[i].first, min_diff);
                step = std::max(min_diff, step);
            }
    }
    if (threshold == 0) threshold = step;
    min_time -= step;  // one spare grid point on either side
    max_time += step;
    const size_t nslots = (size_t)std::llround((max_time - min_time) / step) + 1;
    const size_t ncam = timepairs.size();

    // per (grid point, camera): signed distance of the closest frame and its index
    std::vector<std::pair<double, size_t>> grid(nslots * ncam, std::make_pair(1.01 * threshold, (size_t)0));
    for (size_t cam = 0; cam < ncam; ++cam)
        for (const auto& tp : timepairs[cam]) {
            const long long slot = std::llround((tp.first - min_time) / step);
            for (int d = -1; d <= 1; ++d) {
                const size_t idx = ncam * (size_t)(slot + d) + cam;
                const double delta = tp.first - min_time - (double)(slot + d) * step;
                // "+ epsilon": of two equally distant frames the earlier one wins
                if (std::abs(delta) + kTimeEpsilon < std::abs(grid[idx].first)) grid[idx] = {delta, tp.second};
            }
        }

    double last_delta = 0;
    for (size_t slot = 1; slot + 1 < nslots; ++slot) {
        std::vector<size_t> idx;
        std::vector<double> cam_time;
        const double ftime = min_time + (double)slot * step;
        double total_delta = 0;
        for (size_t cam = 0; cam < ncam; ++cam) {
            const auto& cell = grid[ncam * slot + cam];
            if (std::abs(cell.first) > threshold + kTimeEpsilon) break;
            idx.push_back(cell.second);
            cam_time.push_back(ftime + cell.first);
            total_delta += std::abs(cell.first);
        }
        if (idx.size() != ncam) continue;
        if (frame_indices.empty() || idx != frame_indices.back()) {
            frame_indices.push_back(idx);
            camera_time.push_back(cam_time);
            time.push_back(ftime);
        } else if (total_delta + kTimeEpsilon < last_delta) {
            time.back() = ftime;  // the same frames sit closer to this grid point than to the previous one
        }
        last_delta = total_delta;
    }
}

bool CompositeFrames::read(size_t k, uint64_t offset, uint64_t count, std::vector<double>& out, std::string& error) const {
    out.assign(count, 0.0);
    for (size_t cam = 0; cam < images.size(); ++cam) {
        const auto& im = images[cam];
        const uint64_t lo = std::max(offset, im.pixel_offset), hi = std::min(offset + count, im.pixel_offset + im.npixel);
        if (lo >= hi) continue;
        if (!im.read(frame_indices[k][cam], lo - im.pixel_offset, hi - lo, out.data() + (lo - offset), error)) return false;
    }
    return true;
}
