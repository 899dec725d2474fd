// h5lite -- a small native reader/writer for the subset of HDF5 that the reference's files use, for machines without
// libhdf5 (this image has neither the library nor h5py). NOT verified against libhdf5: the reader is checked against the
// one third-party HDF5 file on the image (scipy's MATLAB v7.3 test file) and both directions against each other.
//
// Reader: superblock versions 0/1, version-1 object headers (with continuation blocks), old-style groups (symbol-table
//   message -> v1 B-tree -> symbol nodes -> local heap), datasets with compact / contiguous / chunked (v1 chunk B-tree,
//   deflate and shuffle filters) layout, layout message versions 1-3, dataspace versions 1-2, attribute versions 1-3,
//   datatypes: fixed-point (any width up to 8 bytes, signed or unsigned), IEEE float 4/8, fixed-length strings and
//   variable-length strings (global heap). Little-endian data only. This is what h5py / libhdf5 1.8-1.14 write by default
//   (libver "earliest"). Of the libver "latest" structures: superblock versions 2/3, version-2 object headers with
//   continuation blocks, new-style groups whose links are link messages in the header (up to 8 links by default), layout
//   message version 4 with contiguous, compact, single-chunk, implicit and unpaged fixed-array chunk storage (<= 1024
//   chunks), filter pipeline version 2. Groups kept in a fractal heap, extensible-array and version-2 B-tree chunk indexes
//   are refused with a clear message. (Checked against an independent pure-Python writer, tests/h5w.py.)
// Writer: superblock 0, old-style groups (one symbol node per group, up to 64 links), contiguous datasets, version-1
//   attribute messages, fixed-length strings, and EXTENDIBLE datasets: chunked layout (version-3 layout message, version-1
//   chunk B-tree), first dimension unlimited, rows appended after the file has been laid out -- what the reference's
//   solution file is (source/solution.cpp:60-165: created with H5S_UNLIMITED, extended every max_cached_solutions frames).
//   An append writes the new chunks and a fresh chunk index at the end of the file, then updates three fixed-size fields
//   (dataset extent, index address, end-of-file address), so the file is a complete HDF5 file after every flush.
#pragma once

#include <cstdint>
#include <cstdio>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace h5lite {

struct Error : std::runtime_error {
    using std::runtime_error::runtime_error;
};

enum class TypeClass { Integer = 0, Float = 1, String = 3, VlenString = 9, Other = 99 };

struct Datatype {
    TypeClass cls = TypeClass::Other;
    uint32_t size = 0;  // bytes per element as stored
    bool is_signed = false;
};

struct Attribute {
    Datatype type;
    std::vector<uint64_t> dims;  // empty = scalar
    std::vector<uint8_t> raw;    // elements as stored (vlen strings already resolved into `strings`)
    std::vector<std::string> strings;
};

struct Dataset {
    Datatype type;
    std::vector<uint64_t> dims;
    int layout = -1;  // 0 compact, 1 contiguous, 2 chunked
    uint64_t address = 0, storage_size = 0;
    std::vector<uint8_t> compact;
    std::vector<uint64_t> chunk_dims;  // without the trailing element-size entry
    std::vector<int> filters;          // 1 deflate, 2 shuffle
    int chunk_index = 0;               // 0: version-1 B-tree; layout version 4: 1 single chunk, 2 implicit, 3 fixed array
    uint64_t single_chunk_size = 0;    // single filtered chunk: its stored size and filter mask
    uint32_t single_filter_mask = 0;
    uint64_t elements() const {
        uint64_t n = 1;
        for (uint64_t d : dims) n *= d;
        return n;
    }
};

struct Object {
    bool is_group = false, is_dataset = false;
    std::map<std::string, uint64_t> links;  // child name -> object header address
    std::map<std::string, Attribute> attributes;
    Dataset dataset;
};

class File {
public:
    explicit File(const std::string& path);
    ~File();
    File(const File&) = delete;
    File& operator=(const File&) = delete;

    bool exists(const std::string& path) const;       // "rtm/voxel_map/i"
    Object object(const std::string& path) const;      // throws Error when missing
    std::vector<std::string> list(const std::string& group_path) const;

    // Whole dataset converted to T (integer <-> integer, float <-> float/double, integer -> floating).
    template <typename T>
    std::vector<T> read(const std::string& path) const;
    // Rows [row0, row0 + nrows) of a dataset whose first dimension is the row index, converted to T.
    template <typename T>
    std::vector<T> read_rows(const std::string& path, uint64_t row0, uint64_t nrows) const;
    // Byte offset of a contiguous, unfiltered dataset's first element in the file (what H5Dget_offset returns), or
    // UINT64_MAX when the dataset is stored otherwise.
    uint64_t contiguous_offset(const std::string& path) const;

    template <typename T>
    T attribute(const std::string& object_path, const std::string& name) const;
    std::string string_attribute(const std::string& object_path, const std::string& name) const;
    bool has_attribute(const std::string& object_path, const std::string& name) const;

    const std::string& path() const { return filename; }

private:
    std::string filename;
    FILE* fp = nullptr;
    uint64_t base = 0, root_header = 0;

    void read_at(uint64_t address, void* dst, size_t bytes) const;
    std::vector<uint8_t> bytes_at(uint64_t address, size_t bytes) const;
    Object parse_object(uint64_t header_address) const;
    Object parse_object_v2(uint64_t header_address) const;
    void parse_messages(const std::vector<uint8_t>& block, Object& obj, std::vector<std::pair<uint64_t, uint64_t>>& more) const;
    void walk_group_btree(uint64_t btree, uint64_t heap_data, std::map<std::string, uint64_t>& links) const;
    void read_raw(const Dataset& d, uint64_t first_element, uint64_t count, std::vector<uint8_t>& out) const;
    void read_chunked(const Dataset& d, std::vector<uint8_t>& whole) const;
    std::string global_heap_string(uint64_t collection, uint32_t index) const;
    uint64_t resolve(const std::string& path) const;
    Attribute parse_attribute(const uint8_t* p, size_t n, std::string& name) const;
};

// ---- writer -------------------------------------------------------------------------------------------------------------
class Writer {
public:
    explicit Writer(const std::string& path);
    ~Writer();
    void create_group(const std::string& path);  // parents must exist; "/" exists
    void write_dataset(const std::string& path, const std::vector<uint64_t>& dims, const double* data);
    void write_dataset(const std::string& path, const std::vector<uint64_t>& dims, const float* data);
    void write_dataset(const std::string& path, const std::vector<uint64_t>& dims, const int32_t* data);
    void write_dataset(const std::string& path, const std::vector<uint64_t>& dims, const uint64_t* data);
    void set_attribute(const std::string& object_path, const std::string& name, double value);
    void set_attribute(const std::string& object_path, const std::string& name, uint64_t value);
    void set_attribute(const std::string& object_path, const std::string& name, int32_t value);
    void set_attribute(const std::string& object_path, const std::string& name, const std::string& value);
    // Extendible dataset of doubles (is_double) or 32-bit signed integers: dims {0, row_shape...}, maximum dims
    // {unlimited, row_shape...}, chunks of chunk_rows rows. Rows are added with append_rows() -- before close() they are
    // buffered and written with the file, after close() they go to the end of the file; flush() makes them part of the
    // dataset (new chunk index, new extent). Reference: Solution::create_hdf5 / update_hdf5, source/solution.cpp:60-165.
    void create_extendible(const std::string& path, bool is_double, const std::vector<uint64_t>& row_shape, uint64_t chunk_rows);
    void append_rows(const std::string& path, const void* rows, uint64_t nrows);
    void flush();
    void close();  // lays the file out and writes it (extendible datasets stay open for append_rows / flush)

private:
    struct Node {
        bool is_group = true;
        std::map<std::string, Node*> children;
        std::vector<std::pair<std::string, std::vector<uint8_t>>> attribute_messages;
        std::vector<uint8_t> datatype, dataspace;  // datasets
        const void* data = nullptr;
        std::vector<uint8_t> owned;
        uint64_t data_bytes = 0;
        uint64_t header_address = 0, btree = 0, heap = 0, heap_data = 0, snod = 0, data_address = 0;
        // extendible datasets
        bool extendible = false;
        std::vector<uint64_t> row_shape;
        uint64_t chunk_rows = 0, row_bytes = 0, rows = 0;         // rows = current extent of the first dimension
        uint32_t elem_size = 0;
        std::vector<uint8_t> pending;                              // rows appended before close()
        std::vector<uint64_t> chunk_address;                       // chunk k holds rows [k * chunk_rows, (k + 1) * chunk_rows)
        uint64_t extent_field = 0, index_field = 0;                // file offsets of dims[0] and of the B-tree address
        bool dirty = false;
        ~Node();
    };
    FILE* append_fp = nullptr;
    uint64_t eof = 0;
    std::vector<Node*> extendibles;
    void write_at(uint64_t address, const void* data, size_t bytes);
    uint64_t write_chunk_index(Node* n);
    void append_to(Node* n, const uint8_t* rows, uint64_t nrows);
    std::string filename;
    Node root;
    bool closed = false;
    Node* find(const std::string& path, bool parent_only, std::string* leaf);
    void add_dataset(const std::string& path, const std::vector<uint64_t>& dims, const std::vector<uint8_t>& datatype,
                     const void* data, uint64_t bytes);
    void add_attribute(const std::string& object_path, const std::string& name, const std::vector<uint8_t>& datatype,
                       const void* data, uint32_t bytes);
};

}  // namespace h5lite
