// Lists an HDF5 file read with h5lite as JSON lines (one per object / attribute), optionally with dataset values:
//   sart_b200_h5ls file.h5 [--values]
// and, for the writer's tests, creates a file that exercises every writer feature:  sart_b200_h5ls --selftest out.h5
#include <cstring>
#include <iostream>

#include "h5lite.hpp"

namespace {

const char* cls_name(h5lite::TypeClass c) {
    switch (c) {
        case h5lite::TypeClass::Integer: return "integer";
        case h5lite::TypeClass::Float: return "float";
        case h5lite::TypeClass::String: return "string";
        case h5lite::TypeClass::VlenString: return "vlen_string";
        default: return "other";
    }
}

void walk(const h5lite::File& f, const std::string& path, bool values) {
    const h5lite::Object o = f.object(path);
    std::cout << "{\"path\": \"" << (path.empty() ? "/" : path) << "\", \"kind\": \"" << (o.is_group ? "group" : o.is_dataset ? "dataset" : "other") << "\"";
    if (o.is_dataset) {
        std::cout << ", \"class\": \"" << cls_name(o.dataset.type.cls) << "\", \"size\": " << o.dataset.type.size
                  << ", \"signed\": " << (o.dataset.type.is_signed ? 1 : 0) << ", \"layout\": " << o.dataset.layout << ", \"dims\": [";
        for (size_t i = 0; i < o.dataset.dims.size(); ++i) std::cout << (i ? ", " : "") << o.dataset.dims[i];
        std::cout << "], \"filters\": [";
        for (size_t i = 0; i < o.dataset.filters.size(); ++i) std::cout << (i ? ", " : "") << o.dataset.filters[i];
        std::cout << "], \"contiguous_offset\": " << (long long)f.contiguous_offset(path);
        if (values && (o.dataset.type.cls == h5lite::TypeClass::Float || o.dataset.type.cls == h5lite::TypeClass::Integer)) {
            const auto v = f.read<double>(path);
            std::cout.precision(17);
            std::cout << ", \"values\": [";
            for (size_t i = 0; i < v.size(); ++i) std::cout << (i ? ", " : "") << v[i];
            std::cout << "]";
        }
    }
    std::cout << ", \"attributes\": {";
    bool first = true;
    for (const auto& kv : o.attributes) {
        std::cout << (first ? "" : ", ") << "\"" << kv.first << "\": ";
        first = false;
        const auto& a = kv.second;
        if (!a.strings.empty()) std::cout << "\"" << a.strings[0] << "\"";
        else if (a.type.cls == h5lite::TypeClass::Float) std::cout << f.attribute<double>(path, kv.first);
        else if (a.type.cls == h5lite::TypeClass::Integer) std::cout << f.attribute<int64_t>(path, kv.first);
        else std::cout << "null";
    }
    std::cout << "}}" << std::endl;
    for (const auto& kv : o.links) walk(f, path.empty() ? kv.first : path + "/" + kv.first, values);
}

}  // namespace

int main(int argc, char* argv[]) {
    try {
        if (argc == 3 && std::strcmp(argv[1], "--selftest") == 0) {
            h5lite::Writer w(argv[2]);
            w.create_group("solution");
            w.create_group("voxel_map");
            w.create_group("solution/nested");
            const double value[6] = {0.5, 1.5, 2.5, 3.5, 4.5, -5.5};
            const double time[2] = {10.0, 20.0};
            const int32_t status[2] = {0, -1};
            const uint64_t idx[3] = {1, 2, 1ull << 40};
            const float small[4] = {1.f, 2.f, 3.f, 4.f};
            w.write_dataset("solution/value", {2, 3}, value);
            w.write_dataset("solution/time", {2}, time);
            w.write_dataset("solution/status", {2}, status);
            w.write_dataset("voxel_map/i", {3}, idx);
            w.write_dataset("solution/nested/small", {2, 2}, small);
            w.set_attribute("voxel_map", "nx", (uint64_t)7);
            w.set_attribute("voxel_map", "xmin", -1.25);
            w.set_attribute("voxel_map", "coordinate_system", std::string("cylindrical"));
            w.set_attribute("solution/value", "flag", (int32_t)-3);
            w.close();
            return 0;
        }
        if (argc < 2) {
            std::cerr << "usage: sart_b200_h5ls file.h5 [--values] | --selftest out.h5" << std::endl;
            return 2;
        }
        h5lite::File f(argv[1]);
        walk(f, "", argc > 2 && std::strcmp(argv[2], "--values") == 0);
    } catch (const std::exception& err) {
        std::cerr << err.what() << std::endl;
        return 1;
    }
    return 0;
}
