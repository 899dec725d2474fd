// Prints the composite frames the time logic selects for the given raw image files and time range, as JSON lines
// (tests/test_frame_source.py): {"time": t, "indices": [...], "camera_time": [...]}
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <iostream>

#include "arguments_b200.hpp"
#include "frame_source.hpp"

int main(int argc, char* argv[]) {
    if (argc < 3) {
        std::cerr << "usage: sart_b200_timesync <time_range or ''> image.img..." << std::endl;
        return 2;
    }
    std::vector<std::array<double, 4>> intervals;
    std::string error;
    if (!parse_time_intervals(argv[1], intervals, error)) {
        std::cerr << error << std::endl;
        return 1;
    }
    std::vector<CameraSeries> images;
    for (int i = 2; i < argc; ++i) {
        rawfiles::ImgHeader h;
        if (!rawfiles::read_img_header(argv[i], h)) {
            std::cerr << "Unable to read " << argv[i] << "." << std::endl;
            return 1;
        }
        images.push_back(camera_from_raw(h));
    }
    try {
        CompositeFrames frames(images, intervals);
        std::cout.precision(12);
        for (size_t k = 0; k < frames.size(); ++k) {
            std::cout << "{\"time\": " << frames.frame_time(k) << ", \"indices\": [";
            for (size_t c = 0; c < frames.indices(k).size(); ++c) std::cout << (c ? ", " : "") << frames.indices(k)[c];
            std::cout << "], \"camera_time\": [";
            for (size_t c = 0; c < frames.camera_frame_time(k).size(); ++c) std::cout << (c ? ", " : "") << frames.camera_frame_time(k)[c];
            std::cout << "]}" << std::endl;
        }
        // SART_TIMESYNC_CHUNK=n: read the pixel window SART_TIMESYNC_WINDOW=first:count (default: everything) of all composite
        // frames in chunks of n (read_chunk) and frame by frame (read); report whether they agree and the reads each took
        if (const char* cs = std::getenv("SART_TIMESYNC_CHUNK")) {
            const size_t chunk = std::max(1, std::atoi(cs));
            uint64_t first = 0, count = frames.npixel();
            if (const char* ws = std::getenv("SART_TIMESYNC_WINDOW")) std::sscanf(ws, "%lu:%lu", &first, &count);
            size_t chunk_reads = 0, frame_reads = 0;
            bool equal = true;
            double checksum = 0;
            std::string err;
            for (size_t k0 = 0; k0 < frames.size(); k0 += chunk) {
                const size_t n = std::min(chunk, frames.size() - k0);
                std::vector<double> all, one;
                size_t r = 0;
                if (!frames.read_chunk(k0, n, first, count, all, err, &r)) throw std::runtime_error(err);
                chunk_reads += r;
                for (size_t b = 0; b < n; ++b) {
                    if (!frames.read(k0 + b, first, count, one, err)) throw std::runtime_error(err);
                    frame_reads += frames.cameras().size();
                    for (uint64_t q = 0; q < count; ++q) {
                        equal = equal && one[q] == all[b * count + q];
                        checksum += one[q] * (double)(q + 1);
                    }
                }
            }
            std::cout << "{\"chunk_reads\": " << chunk_reads << ", \"frame_reads\": " << frame_reads << ", \"equal\": " << (equal ? 1 : 0)
                      << ", \"checksum\": " << checksum << "}" << std::endl;
        }
    } catch (const std::runtime_error& err) {
        std::cerr << err.what() << std::endl;
        return 1;
    }
    return 0;
}
