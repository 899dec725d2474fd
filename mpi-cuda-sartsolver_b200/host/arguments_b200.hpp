// Command line of the reference's `sartsolver`, re-implemented without the argparse dependency: the same 17 options
// with the same names, short forms, defaults and range checks (source/arguments.cpp:86-236; `Config`,
// source/include/arguments.hpp:14-33) and the same time-range mini-language (source/arguments.cpp:12-79).
// Parsing never exits: it returns false with the message the reference would have printed, so that it can be tested;
// parse_arguments_or_exit() adds the reference's behaviour (message on std::cerr, exit(1)).
#pragma once

#include <array>
#include <string>
#include <vector>

struct Config {
    std::vector<std::string> input_files;
    std::string output_file = "solution.h5";
    std::string time_range;
    std::string laplacian_file;
    std::string raytransfer_name = "with_reflections";
    double wavelength_threshold = 50.;
    double ray_density_threshold = 1.e-6;
    double ray_length_threshold = 1.e-6;
    double conv_tolerance = 1.e-5;
    double beta_laplace = 2.e-2;
    double relaxation = 1.;
    int max_iterations = 2000;
    int max_cached_frames = 100;
    int max_cached_solutions = 100;
    bool logarithmic = false;
    bool no_guess = false;
    bool use_cpu = false;
    bool parallel_read = false;
};

bool parse_arguments(int argc, const char* const argv[], Config& config, std::string& error);
Config parse_arguments_or_exit(int argc, char* argv[]);

// "start:stop[:step[:sync_threshold]]", comma separated; empty -> one interval [0, inf). Each entry is
// {start, stop, step, threshold}.
bool parse_time_intervals(const std::string& time_string, std::vector<std::array<double, 4>>& intervals,
                          std::string& error);

std::string usage();
