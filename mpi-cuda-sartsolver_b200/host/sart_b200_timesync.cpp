// Prints the composite frames the time logic selects for the given raw image files and time range, as JSON lines
// (tests/test_frame_source.py): {"time": t, "indices": [...], "camera_time": [...]}
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
    } catch (const std::runtime_error& err) {
        std::cerr << err.what() << std::endl;
        return 1;
    }
    return 0;
}
