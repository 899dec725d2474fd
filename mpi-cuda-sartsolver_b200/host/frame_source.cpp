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
    // consecutive frames are consecutive in the file: whole-camera reads of a run are one contiguous read
    c.read_run = [path, data_offset, npixel](size_t index, size_t nframes, uint64_t first, uint64_t count, double* out,
                                             uint64_t out_stride, std::string& error) {
        FILE* f = std::fopen(path.c_str(), "rb");
        if (!f) {
            error = "cannot open " + path;
            return false;
        }
        bool ok = true;
        if (count == npixel && out_stride == npixel) {
            ok = fseeko(f, (off_t)(data_offset + index * npixel * 8), SEEK_SET) == 0 && rawfiles::read_exact(f, out, nframes * npixel * 8);
        } else {
            for (size_t k = 0; k < nframes && ok; ++k)
                ok = fseeko(f, (off_t)(data_offset + ((index + k) * npixel + first) * 8), SEEK_SET) == 0 &&
                     rawfiles::read_exact(f, out + k * out_stride, count * 8);
        }
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

// Multi-camera synchronisation (the behaviour of source/image.cpp:110-196, written from its contract, as a gather):
// a time grid of pitch `step` is laid over the interval, starting one step before the earliest frame of any camera and
// ending one step after the latest; a composite frame exists at a grid point when EVERY camera has a frame within
// `threshold` of it. step 0 = the largest of the cameras' smallest frame spacings; threshold 0 = step.
namespace {

using TimedFrames = std::vector<std::pair<double, size_t>>;  // one camera: (time, frame index), ascending in time

struct TimeGrid {
    double origin = 0, step = 1, threshold = 1;
    size_t npoints = 0;
    double at(size_t k) const { return origin + (double)k * step; }
    // the grid point a frame belongs to: the nearest one
    long long home(double t) const { return std::llround((t - origin) / step); }
};

// The smallest spacing between two consecutive frames of a camera (its whole span when it has a single spacing).
double finest_spacing(const TimedFrames& frames) {
    double finest = frames.back().first - frames.front().first;
    for (size_t i = 1; i < frames.size(); ++i) finest = std::min(finest, frames[i].first - frames[i - 1].first);
    return finest;
}

struct NearestFrame {
    double offset;  // frame time - grid time
    size_t frame;
};

// Of the frames whose home is grid point k or one of its two neighbours, the one closest to k. Walking them in time
// order, a later frame only displaces an earlier one when it is closer by more than kTimeEpsilon (equidistant frames:
// the earlier one is kept), and nothing further than 1 % beyond the threshold is ever considered.
NearestFrame nearest_frame(const TimedFrames& frames, const TimeGrid& grid, size_t k) {
    NearestFrame best{1.01 * grid.threshold, 0};
    const long long lo = (long long)k - 1, hi = (long long)k + 1;
    auto it = std::partition_point(frames.begin(), frames.end(), [&](const auto& f) { return grid.home(f.first) < lo; });
    for (; it != frames.end() && grid.home(it->first) <= hi; ++it) {
        const double offset = it->first - grid.at(k);
        if (std::abs(offset) + kTimeEpsilon < std::abs(best.offset)) best = {offset, it->second};
    }
    return best;
}

}  // namespace

void CompositeFrames::add_interval(const std::vector<std::vector<std::pair<double, size_t>>>& timepairs, double step,
                                   double threshold) {
    const size_t ncam = timepairs.size();
    if (ncam == 0) return;
    double earliest = 0, latest = 0;
    for (size_t cam = 0; cam < ncam; ++cam) {
        if (timepairs[cam].empty()) return;  // a camera without frames in the interval: nothing can be combined
        const double first = timepairs[cam].front().first, last = timepairs[cam].back().first;
        earliest = cam ? std::min(earliest, first) : first;
        latest = cam ? std::max(latest, last) : last;
    }
    TimeGrid grid;
    if (step != 0) grid.step = step;
    else if (latest - earliest < kTimeEpsilon) grid.step = 1.;  // every frame at the same instant
    else {
        grid.step = 0;
        for (const auto& frames : timepairs) grid.step = std::max(grid.step, finest_spacing(frames));
    }
    grid.threshold = threshold != 0 ? threshold : grid.step;
    grid.origin = earliest - grid.step;
    grid.npoints = (size_t)std::llround((latest + grid.step - grid.origin) / grid.step) + 1;

    std::vector<size_t> chosen(ncam);
    std::vector<double> chosen_time(ncam);
    double previous_misfit = 0;  // summed |offset| of the last grid point that had a frame of every camera
    for (size_t k = 1; k + 1 < grid.npoints; ++k) {  // the first and the last point are the spare ones
        double misfit = 0;
        size_t cam = 0;
        for (; cam < ncam; ++cam) {
            const NearestFrame pick = nearest_frame(timepairs[cam], grid, k);
            if (std::abs(pick.offset) > grid.threshold + kTimeEpsilon) break;
            chosen[cam] = pick.frame;
            chosen_time[cam] = grid.at(k) + pick.offset;
            misfit += std::abs(pick.offset);
        }
        if (cam < ncam) continue;  // some camera has nothing close enough
        if (frame_indices.empty() || chosen != frame_indices.back()) {
            frame_indices.push_back(chosen);
            camera_time.push_back(chosen_time);
            time.push_back(grid.at(k));
        } else if (misfit + kTimeEpsilon < previous_misfit) {
            time.back() = grid.at(k);  // the same frames again, and this grid point fits them better than the one before
        }
        previous_misfit = misfit;
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

bool CompositeFrames::read_chunk(size_t k0, size_t n, uint64_t offset, uint64_t count, std::vector<double>& out, std::string& error,
                                 size_t* reads) const {
    out.assign(n * count, 0.0);
    size_t issued = 0;
    for (size_t cam = 0; cam < images.size(); ++cam) {
        const auto& im = images[cam];
        const uint64_t lo = std::max(offset, im.pixel_offset), hi = std::min(offset + count, im.pixel_offset + im.npixel);
        if (lo >= hi) continue;
        double* dst = out.data() + (lo - offset);
        size_t b = 0;
        while (b < n) {
            // the longest run of composite frames whose frames of this camera follow each other in its file
            size_t e = b + 1;
            while (e < n && frame_indices[k0 + e][cam] == frame_indices[k0 + e - 1][cam] + 1) ++e;
            const size_t first_index = frame_indices[k0 + b][cam];
            if (im.read_run) {
                if (!im.read_run(first_index, e - b, lo - im.pixel_offset, hi - lo, dst + b * count, count, error)) return false;
                ++issued;
            } else {
                for (size_t k = b; k < e; ++k) {
                    if (!im.read(frame_indices[k0 + k][cam], lo - im.pixel_offset, hi - lo, dst + k * count, error)) return false;
                    ++issued;
                }
            }
            b = e;
        }
    }
    if (reads) *reads = issued;
    return true;
}
