#include "arguments_b200.hpp"

#include <cstdlib>
#include <iostream>
#include <limits>
#include <map>
#include <sstream>

namespace {

struct Option {
    const char* short_name;
    const char* long_name;
    bool takes_value;
    const char* help;
};

const Option kOptions[] = {
    {"-o", "--output_file", true, "Filename to save the solution."},
    {"-t", "--time_range", true,
     "Time intervals in s to process in a form: start:stop:(step):(synch_threshold), e.g. '20.5:40.1, 45.2:51:15:0.05'."},
    {"-w", "--wavelength_threshold", true,
     "An RTM is considered valid if its wavelength is within this threshold of the image wavelength (in nm)."},
    {"-d", "--ray_density_threshold", true, "Voxels with ray density lesser than this threshold are ignored."},
    {"-r", "--ray_length_threshold", true, "Pixels with ray length lesser than this threshold are ignored."},
    {"-m", "--max_iterations", true, "Maximum number of SART iterations."},
    {"-c", "--conv_tolerance", true, "SART convolution relative tolerance."},
    {"-l", "--laplacian_file", true, "File with laplacian regularization matrix."},
    {"-b", "--beta_laplace", true, "Weight of the regularization factor."},
    {"-R", "--relaxation", true, "Relaxation parameter."},
    {"-n", "--raytransfer_name", true, "Ray transfer matrix dataset name."},
    {"-L", "--logarithmic", false, "Use logarithmic SART solver."},
    {nullptr, "--max_cached_frames", true, "Maximum number of cached image frames."},
    {nullptr, "--max_cached_solutions", true, "Maximum number of cached solutions."},
    {nullptr, "--no_guess", false,
     "Do not use solution found on previous time moment as initial guess for the next one."},
    {nullptr, "--use_cpu", false, "Perform all calculations on CPUs."},
    {nullptr, "--parallel_read", false, "Read RTM data in a parallel way (high-IOPS storage optimization)."},
};

const Option* find_option(const std::string& name) {
    for (const Option& o : kOptions)
        if (name == o.long_name || (o.short_name && name == o.short_name)) return &o;
    return nullptr;
}

bool to_double(const std::string& text, const char* what, double& out, std::string& error) {
    char* end = nullptr;
    out = std::strtod(text.c_str(), &end);
    if (text.empty() || end != text.c_str() + text.size()) {
        error = std::string("Failed to parse '") + text + "' as a number for " + what + ".";
        return false;
    }
    return true;
}

bool to_int(const std::string& text, const char* what, int& out, std::string& error) {
    char* end = nullptr;
    const long v = std::strtol(text.c_str(), &end, 10);
    if (text.empty() || end != text.c_str() + text.size() || v < std::numeric_limits<int>::min() ||
        v > std::numeric_limits<int>::max()) {
        error = std::string("Failed to parse '") + text + "' as an integer for " + what + ".";
        return false;
    }
    out = (int)v;
    return true;
}

template <typename T>
std::string shown(T v) {
    std::ostringstream os;
    os << v;
    return os.str();
}

}  // namespace

std::string usage() {
    std::ostringstream os;
    os << "Usage: sartsolver [options] input_files...\n\nPositional arguments:\n"
       << "  input_files           List of ray transfer matrix and camera image files.\n\nOptional arguments:\n";
    for (const Option& o : kOptions) {
        std::string names = o.short_name ? std::string(o.short_name) + ", " + o.long_name : std::string(o.long_name);
        os << "  " << names << (o.takes_value ? " VALUE" : "") << "\n        " << o.help << "\n";
    }
    return os.str();
}

bool parse_arguments(int argc, const char* const argv[], Config& config, std::string& error) {
    config = Config();
    std::map<std::string, std::string> values;
    int i = 1;
    for (; i < argc; ++i) {
        std::string arg = argv[i];
        if (arg.empty() || arg[0] != '-' || arg == "-") break;  // first positional: everything from here is a file
        if (arg == "--") {
            ++i;
            break;
        }
        std::string inline_value;
        bool has_inline = false;
        const size_t eq = arg.find('=');
        if (arg.rfind("--", 0) == 0 && eq != std::string::npos) {
            inline_value = arg.substr(eq + 1);
            arg = arg.substr(0, eq);
            has_inline = true;
        }
        const Option* opt = find_option(arg);
        if (!opt) {
            // a negative number cannot be an option: it starts the positionals only if nothing claims it
            error = "Unknown argument: " + arg;
            return false;
        }
        if (!opt->takes_value) {
            if (has_inline) {
                error = std::string("Argument ") + opt->long_name + " takes no value.";
                return false;
            }
            values[opt->long_name] = "1";
            continue;
        }
        if (has_inline) values[opt->long_name] = inline_value;
        else {
            if (i + 1 >= argc) {
                error = std::string("Too few arguments for '") + arg + "'.";
                return false;
            }
            values[opt->long_name] = argv[++i];
        }
    }
    for (; i < argc; ++i) config.input_files.push_back(argv[i]);

    auto has = [&](const char* name) { return values.count(name) != 0; };
    // same order of checks and wording as source/arguments.cpp:184-236
    if (has("--ray_density_threshold") &&
        !to_double(values["--ray_density_threshold"], "ray_density_threshold", config.ray_density_threshold, error))
        return false;
    if (config.ray_density_threshold < 0) {
        error = "Argument ray_density_threshold must be >= 0, " + shown(config.ray_density_threshold) + " given.";
        return false;
    }
    if (has("--ray_length_threshold") &&
        !to_double(values["--ray_length_threshold"], "ray_length_threshold", config.ray_length_threshold, error))
        return false;
    if (config.ray_length_threshold < 0) {
        error = "Argument ray_length_threshold must be >= 0, " + shown(config.ray_length_threshold) + " given.";
        return false;
    }
    if (has("--max_iterations") && !to_int(values["--max_iterations"], "max_iterations", config.max_iterations, error))
        return false;
    if (config.max_iterations < 1) {
        error = "Argument max_iterations must be >= 1, " + shown(config.max_iterations) + " given.";
        return false;
    }
    if (has("--conv_tolerance") && !to_double(values["--conv_tolerance"], "conv_tolerance", config.conv_tolerance, error))
        return false;
    if (!(config.conv_tolerance > 0)) {
        error = "Argument conv_tolerance must be > 0, " + shown(config.conv_tolerance) + " given.";
        return false;
    }
    if (has("--relaxation") && !to_double(values["--relaxation"], "relaxation", config.relaxation, error)) return false;
    if (!(config.relaxation > 0) || config.relaxation > 1.) {
        error = "Argument relaxation must be within (0, 1] interval," + shown(config.relaxation) + " given.";
        return false;
    }
    if (has("--beta_laplace") && !to_double(values["--beta_laplace"], "beta_laplace", config.beta_laplace, error)) return false;
    if (config.beta_laplace < 0) {
        error = "Argument beta_laplace must be positive.";
        return false;
    }
    if (has("--max_cached_frames") &&
        !to_int(values["--max_cached_frames"], "max_cached_frames", config.max_cached_frames, error))
        return false;
    if (config.max_cached_frames <= 0) {
        error = "Argument max_cached_frames must be positive.";
        return false;
    }
    if (has("--max_cached_solutions") &&
        !to_int(values["--max_cached_solutions"], "max_cached_solutions", config.max_cached_solutions, error))
        return false;
    if (config.max_cached_solutions <= 0) {
        error = "Argument max_cached_solutions must be positive.";
        return false;
    }
    if (config.input_files.size() < 2) {
        error = "At least two input file, one with RTM and one with image, are required, " +
                shown(config.input_files.size()) + " given.";
        return false;
    }
    if (has("--wavelength_threshold") &&
        !to_double(values["--wavelength_threshold"], "wavelength_threshold", config.wavelength_threshold, error))
        return false;
    if (has("--output_file")) config.output_file = values["--output_file"];
    if (has("--time_range")) config.time_range = values["--time_range"];
    if (has("--raytransfer_name")) config.raytransfer_name = values["--raytransfer_name"];
    if (has("--laplacian_file")) config.laplacian_file = values["--laplacian_file"];
    config.logarithmic = has("--logarithmic");
    config.parallel_read = has("--parallel_read");
    config.use_cpu = has("--use_cpu");
    config.no_guess = has("--no_guess");
    return true;
}

Config parse_arguments_or_exit(int argc, char* argv[]) {
    Config config;
    std::string error;
    if (!parse_arguments(argc, argv, config, error)) {
        std::cerr << error << std::endl << usage();
        std::exit(1);
    }
    return config;
}

bool parse_time_intervals(const std::string& time_string, std::vector<std::array<double, 4>>& intervals,
                          std::string& error) {
    intervals.clear();
    if (time_string.empty()) {
        intervals.push_back({0, std::numeric_limits<double>::infinity(), 0, 0});
        return true;
    }
    // split on ','; one trailing ',' is allowed (the reference stops one character before the end, :25)
    size_t pos = 0;
    while (pos + 1 < time_string.size()) {
        const size_t comma = time_string.find(',', pos);
        const std::string item = time_string.substr(pos, comma == std::string::npos ? std::string::npos : comma - pos);
        pos = (comma == std::string::npos) ? time_string.size() : comma + 1;

        std::vector<std::string> fields;
        size_t fpos = 0;
        while (fpos < item.size()) {
            const size_t colon = item.find(':', fpos);
            fields.push_back(item.substr(fpos, colon == std::string::npos ? std::string::npos : colon - fpos));
            fpos = (colon == std::string::npos) ? item.size() : colon + 1;
        }
        if (fields.size() < 2) {
            error = "Unable to recognize a time interval in " + item + ".";
            return false;
        }
        if (fields.size() > 4) {
            error = "Too many values in a time interval: " + item + ".";
            return false;
        }
        double v[4] = {0, 0, 0, 0};
        for (size_t k = 0; k < fields.size(); ++k) {
            // std::stod semantics: leading blanks are skipped, trailing characters ignored, nothing parsed = error
            char* end = nullptr;
            v[k] = std::strtod(fields[k].c_str(), &end);
            if (end == fields[k].c_str()) {
                error = "Unable to convert " + item + " to the time interval.";
                return false;
            }
        }
        if (v[0] < 0) {
            error = "Time limits must be positive.";
            return false;
        }
        if (v[1] <= v[0]) {
            error = "The upper limit of the time interval must be higher than the lower one.";
            return false;
        }
        if (v[2] > v[1] - v[0]) {
            error = "Time step must be less or equal to the time interval.";
            return false;
        }
        if (v[3] > v[2]) {
            error = "Synchronization threshold must be less or equal to the time step.";
            return false;
        }
        intervals.push_back({v[0], v[1], v[2], v[3]});
    }
    return true;
}
