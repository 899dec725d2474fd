// Prints what the command line parses to, as JSON (exit 0), or the error (exit 1). Used by tests/test_cli.py.
#include <cmath>
#include <iostream>

#include "arguments_b200.hpp"

int main(int argc, char* argv[]) {
    Config c;
    std::string error;
    if (!parse_arguments(argc, argv, c, error)) {
        std::cerr << error << std::endl;
        return 1;
    }
    std::vector<std::array<double, 4>> iv;
    if (!parse_time_intervals(c.time_range, iv, error)) {
        std::cerr << error << std::endl;
        return 1;
    }
    std::cout.precision(17);
    std::cout << "{\"input_files\": [";
    for (size_t i = 0; i < c.input_files.size(); ++i) std::cout << (i ? ", " : "") << '"' << c.input_files[i] << '"';
    std::cout << "], \"output_file\": \"" << c.output_file << "\", \"time_range\": \"" << c.time_range
              << "\", \"laplacian_file\": \"" << c.laplacian_file << "\", \"raytransfer_name\": \"" << c.raytransfer_name
              << "\", \"wavelength_threshold\": " << c.wavelength_threshold
              << ", \"ray_density_threshold\": " << c.ray_density_threshold
              << ", \"ray_length_threshold\": " << c.ray_length_threshold << ", \"conv_tolerance\": " << c.conv_tolerance
              << ", \"beta_laplace\": " << c.beta_laplace << ", \"relaxation\": " << c.relaxation
              << ", \"max_iterations\": " << c.max_iterations << ", \"max_cached_frames\": " << c.max_cached_frames
              << ", \"max_cached_solutions\": " << c.max_cached_solutions << ", \"logarithmic\": " << c.logarithmic
              << ", \"no_guess\": " << c.no_guess << ", \"use_cpu\": " << c.use_cpu
              << ", \"parallel_read\": " << c.parallel_read << ", \"time_intervals\": [";
    for (size_t i = 0; i < iv.size(); ++i) {
        std::cout << (i ? ", " : "") << "[";
        for (int k = 0; k < 4; ++k) {
            if (k) std::cout << ", ";
            if (std::isinf(iv[i][k])) std::cout << "\"inf\"";
            else std::cout << iv[i][k];
        }
        std::cout << "]";
    }
    std::cout << "]}" << std::endl;
    return 0;
}
