// C entry points over h5lite::Writer so that the Python tests can compose HDF5 files in the reference's schemas
// (libh5lite.so; test tooling, not part of the solver library).
#include <cstring>
#include <string>
#include <vector>

#include "h5lite.hpp"

namespace {
thread_local std::string g_err;
template <typename F>
int guard(F f) {
    try {
        f();
        return 0;
    } catch (const std::exception& e) {
        g_err = e.what();
        return 1;
    }
}
std::vector<uint64_t> dims_of(const uint64_t* dims, int rank) { return std::vector<uint64_t>(dims, dims + rank); }
}  // namespace

extern "C" {
const char* h5w_error() { return g_err.c_str(); }
void* h5w_open(const char* path) {
    try {
        return new h5lite::Writer(path);
    } catch (const std::exception& e) {
        g_err = e.what();
        return nullptr;
    }
}
int h5w_group(void* w, const char* path) { return guard([&] { static_cast<h5lite::Writer*>(w)->create_group(path); }); }
int h5w_f64(void* w, const char* path, const uint64_t* dims, int rank, const double* d) {
    return guard([&] { static_cast<h5lite::Writer*>(w)->write_dataset(path, dims_of(dims, rank), d); });
}
int h5w_f32(void* w, const char* path, const uint64_t* dims, int rank, const float* d) {
    return guard([&] { static_cast<h5lite::Writer*>(w)->write_dataset(path, dims_of(dims, rank), d); });
}
int h5w_i32(void* w, const char* path, const uint64_t* dims, int rank, const int32_t* d) {
    return guard([&] { static_cast<h5lite::Writer*>(w)->write_dataset(path, dims_of(dims, rank), d); });
}
int h5w_u64(void* w, const char* path, const uint64_t* dims, int rank, const uint64_t* d) {
    return guard([&] { static_cast<h5lite::Writer*>(w)->write_dataset(path, dims_of(dims, rank), d); });
}
int h5w_attr_f64(void* w, const char* obj, const char* name, double v) {
    return guard([&] { static_cast<h5lite::Writer*>(w)->set_attribute(obj, name, v); });
}
int h5w_attr_u64(void* w, const char* obj, const char* name, uint64_t v) {
    return guard([&] { static_cast<h5lite::Writer*>(w)->set_attribute(obj, name, v); });
}
int h5w_attr_i32(void* w, const char* obj, const char* name, int32_t v) {
    return guard([&] { static_cast<h5lite::Writer*>(w)->set_attribute(obj, name, v); });
}
int h5w_attr_str(void* w, const char* obj, const char* name, const char* v) {
    return guard([&] { static_cast<h5lite::Writer*>(w)->set_attribute(obj, name, std::string(v)); });
}
// extendible datasets (rows appended before or after h5w_layout; h5w_flush makes appended rows part of the dataset)
int h5w_extendible(void* w, const char* path, int is_double, const uint64_t* row_shape, int row_rank, uint64_t chunk_rows) {
    return guard([&] { static_cast<h5lite::Writer*>(w)->create_extendible(path, is_double != 0, dims_of(row_shape, row_rank), chunk_rows); });
}
int h5w_append(void* w, const char* path, const void* rows, uint64_t nrows) {
    return guard([&] { static_cast<h5lite::Writer*>(w)->append_rows(path, rows, nrows); });
}
int h5w_layout(void* w) { return guard([&] { static_cast<h5lite::Writer*>(w)->close(); }); }
int h5w_flush(void* w) { return guard([&] { static_cast<h5lite::Writer*>(w)->flush(); }); }
int h5w_close(void* w) {
    const int rc = guard([&] { static_cast<h5lite::Writer*>(w)->close(); });
    delete static_cast<h5lite::Writer*>(w);
    return rc;
}
}
