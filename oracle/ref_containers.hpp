// TEST INFRASTRUCTURE ONLY -- see ref_containers.cpp. Where the stand-in read_hdf5 members take their data from.
#pragma once
#include <cstddef>

void ref_containers_set_sources(const float* matrix, const size_t* lap_index, const float* lap_value, size_t lap_nnz);
