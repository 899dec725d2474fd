// See h5lite.hpp. Layouts follow the HDF5 File Format Specification, version 1.x structures ("classic" format).
#include "h5lite.hpp"

#include <zlib.h>

#include <algorithm>
#include <cstring>
#include <limits>

namespace h5lite {

namespace {

constexpr uint64_t kUndefined = std::numeric_limits<uint64_t>::max();
const uint8_t kSignature[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};

template <typename T>
T get(const uint8_t* p) {
    T v;
    std::memcpy(&v, p, sizeof(T));
    return v;
}
inline uint64_t pad8(uint64_t n) { return (n + 7) & ~uint64_t(7); }

Datatype parse_datatype(const uint8_t* p, size_t n) {
    if (n < 8) throw Error("truncated datatype message");
    Datatype t;
    const int cls = p[0] & 0x0f;
    const uint8_t bits0 = p[1];
    t.size = get<uint32_t>(p + 4);
    if (cls == 0) {
        if (bits0 & 1) throw Error("big-endian integers are not supported");
        t.cls = TypeClass::Integer;
        t.is_signed = (bits0 & 0x08) != 0;
    } else if (cls == 1) {
        if (bits0 & 1) throw Error("big-endian floats are not supported");
        t.cls = TypeClass::Float;
        if (t.size != 4 && t.size != 8) throw Error("only 4- and 8-byte IEEE floats are supported");
    } else if (cls == 3) {
        t.cls = TypeClass::String;
    } else if (cls == 9) {
        // variable length: class bits 0-3 = type (0 sequence, 1 string)
        t.cls = ((bits0 & 0x0f) == 1) ? TypeClass::VlenString : TypeClass::Other;
    } else {
        t.cls = TypeClass::Other;
    }
    return t;
}

void parse_dataspace(const uint8_t* p, size_t n, std::vector<uint64_t>& dims) {
    dims.clear();
    if (n < 4) throw Error("truncated dataspace message");
    const int version = p[0], rank = p[1];
    size_t off;
    if (version == 1) off = 8;
    else if (version == 2) {
        off = 4;
        if (p[3] == 2) return;  // null dataspace
    } else throw Error("unsupported dataspace message version " + std::to_string(version));
    if (n < off + (size_t)rank * 8) throw Error("truncated dataspace message");
    for (int i = 0; i < rank; ++i) dims.push_back(get<uint64_t>(p + off + 8 * i));
}

template <typename T>
T convert_one(const uint8_t* p, const Datatype& t) {
    if (t.cls == TypeClass::Float) return (t.size == 4) ? (T)get<float>(p) : (T)get<double>(p);
    if (t.cls == TypeClass::Integer) {
        uint64_t u = 0;
        std::memcpy(&u, p, std::min<size_t>(t.size, 8));
        if (t.is_signed && t.size < 8 && (u >> (8 * t.size - 1)) & 1) u |= ~uint64_t(0) << (8 * t.size);
        return t.is_signed ? (T)(int64_t)u : (T)u;
    }
    throw Error("dataset or attribute is not numeric");
}

void unshuffle(std::vector<uint8_t>& buf, size_t elem) {
    if (elem <= 1) return;
    const size_t n = buf.size() / elem;
    std::vector<uint8_t> out(buf.size());
    for (size_t b = 0; b < elem; ++b)
        for (size_t i = 0; i < n; ++i) out[i * elem + b] = buf[b * n + i];
    std::copy(buf.begin() + n * elem, buf.end(), out.begin() + n * elem);
    buf.swap(out);
}

}  // namespace

// ---- reader -------------------------------------------------------------------------------------------------------------

File::File(const std::string& path) : filename(path) {
    fp = std::fopen(path.c_str(), "rb");
    if (!fp) throw Error("cannot open " + path);
    // the superblock may sit at 0, 512, 1024, ... (user block)
    bool found = false;
    uint8_t sig[8];
    for (uint64_t off = 0; off < (uint64_t(1) << 26); off = off ? off * 2 : 512) {
        if (std::fseek(fp, (long)off, SEEK_SET) != 0 || std::fread(sig, 1, 8, fp) != 8) break;
        if (std::memcmp(sig, kSignature, 8) == 0) {
            base = off;
            found = true;
            break;
        }
    }
    if (!found) throw Error(path + " is not an HDF5 file");
    uint8_t sb[96];
    read_at(0, sb, 8 + 48);
    const int version = sb[8];
    if (version == 2 || version == 3) {
        // superblock versions 2 / 3 (libver "latest"): sizes of offsets / lengths, flags, base address, superblock extension,
        // end of file, root group OBJECT HEADER address, checksum
        if (sb[9] != 8 || sb[10] != 8) throw Error(path + ": only 8-byte offsets and lengths are supported");
        const uint64_t base_address = get<uint64_t>(sb + 12);
        if (base_address != 0 && base_address != base) base = base_address;
        root_header = get<uint64_t>(sb + 36);
        return;
    }
    if (version > 3) throw Error(path + ": superblock version " + std::to_string(version) + " is not supported by this reader");
    if (sb[13] != 8 || sb[14] != 8) throw Error(path + ": only 8-byte offsets and lengths are supported");
    // v0: 24 bytes of fixed fields, v1 has 4 more (indexed storage K + reserved)
    const size_t fixed = (version == 0) ? 24 : 28;
    read_at(0, sb, fixed + 32 + 40);
    const uint64_t base_address = get<uint64_t>(sb + fixed);
    if (base_address != 0 && base_address != base) base = base_address;  // absolute base address given
    // root group symbol table entry: link name offset, object header address, cache type, reserved, scratch
    root_header = get<uint64_t>(sb + fixed + 32 + 8);
}

File::~File() {
    if (fp) std::fclose(fp);
}

void File::read_at(uint64_t address, void* dst, size_t bytes) const {
    if (bytes == 0) return;
    if (fseeko(fp, (off_t)(base + address), SEEK_SET) != 0 || std::fread(dst, 1, bytes, fp) != bytes)
        throw Error("short read in " + filename);
}

std::vector<uint8_t> File::bytes_at(uint64_t address, size_t bytes) const {
    std::vector<uint8_t> v(bytes);
    read_at(address, v.data(), bytes);
    return v;
}

std::string File::global_heap_string(uint64_t collection, uint32_t index) const {
    uint8_t head[16];
    read_at(collection, head, 16);
    if (std::memcmp(head, "GCOL", 4) != 0) throw Error("bad global heap collection in " + filename);
    const uint64_t size = get<uint64_t>(head + 8);
    const std::vector<uint8_t> col = bytes_at(collection, size);
    size_t p = 16;
    while (p + 16 <= col.size()) {
        const uint16_t idx = get<uint16_t>(&col[p]);
        const uint64_t len = get<uint64_t>(&col[p + 8]);
        if (idx == 0) break;
        if (idx == index) return std::string(reinterpret_cast<const char*>(&col[p + 16]), len);
        p += 16 + pad8(len);
    }
    throw Error("global heap object not found in " + filename);
}

Attribute File::parse_attribute(const uint8_t* p, size_t n, std::string& name) const {
    if (n < 8) throw Error("truncated attribute message");
    const int version = p[0];
    const uint16_t name_size = get<uint16_t>(p + 2), type_size = get<uint16_t>(p + 4), space_size = get<uint16_t>(p + 6);
    size_t off = (version == 3) ? 9 : 8;
    auto step = [&](uint64_t s) { return version == 1 ? pad8(s) : s; };
    if (version < 1 || version > 3) throw Error("unsupported attribute message version");
    if (off + step(name_size) + step(type_size) + step(space_size) > n) throw Error("truncated attribute message");
    name.assign(reinterpret_cast<const char*>(p + off), name_size);
    name = name.c_str();  // cut at the terminating NUL
    off += step(name_size);
    Attribute a;
    a.type = parse_datatype(p + off, type_size);
    off += step(type_size);
    parse_dataspace(p + off, space_size, a.dims);
    off += step(space_size);
    uint64_t count = 1;
    for (uint64_t d : a.dims) count *= d;
    const uint64_t bytes = count * a.type.size;
    if (off + bytes > n) throw Error("truncated attribute data");
    a.raw.assign(p + off, p + off + bytes);
    if (a.type.cls == TypeClass::String) {
        for (uint64_t i = 0; i < count; ++i) {
            std::string s(reinterpret_cast<const char*>(&a.raw[i * a.type.size]), a.type.size);
            a.strings.push_back(s.c_str());
        }
    } else if (a.type.cls == TypeClass::VlenString) {
        // element = length (4), global heap collection address (8), object index (4)
        for (uint64_t i = 0; i < count; ++i) {
            const uint8_t* e = &a.raw[i * 16];
            a.strings.push_back(global_heap_string(get<uint64_t>(e + 4), get<uint32_t>(e + 12)));
        }
    }
    return a;
}

void File::walk_group_btree(uint64_t btree, uint64_t heap_data, std::map<std::string, uint64_t>& links) const {
    uint8_t head[24];
    read_at(btree, head, 24);
    if (std::memcmp(head, "TREE", 4) != 0 || head[4] != 0) throw Error("bad group B-tree node in " + filename);
    const int level = head[5];
    const int entries = get<uint16_t>(head + 6);
    const std::vector<uint8_t> body = bytes_at(btree + 24, (size_t)entries * 16 + 8);
    for (int i = 0; i < entries; ++i) {
        const uint64_t child = get<uint64_t>(&body[8 + 16 * i]);
        if (level > 0) {
            walk_group_btree(child, heap_data, links);
            continue;
        }
        uint8_t sn[8];
        read_at(child, sn, 8);
        if (std::memcmp(sn, "SNOD", 4) != 0) throw Error("bad symbol table node in " + filename);
        const int nsym = get<uint16_t>(sn + 6);
        const std::vector<uint8_t> ent = bytes_at(child + 8, (size_t)nsym * 40);
        for (int s = 0; s < nsym; ++s) {
            const uint64_t name_off = get<uint64_t>(&ent[40 * s]);
            const uint64_t header = get<uint64_t>(&ent[40 * s + 8]);
            std::string name;
            char c;
            for (uint64_t k = 0;; ++k) {
                read_at(heap_data + name_off + k, &c, 1);
                if (!c) break;
                name.push_back(c);
            }
            links[name] = header;
        }
    }
}

void File::parse_messages(const std::vector<uint8_t>& block, Object& obj, std::vector<std::pair<uint64_t, uint64_t>>& more) const {
    size_t p = 0;
    while (p + 8 <= block.size()) {
        const uint16_t type = get<uint16_t>(&block[p]);
        const uint16_t size = get<uint16_t>(&block[p + 2]);
        const uint8_t* body = &block[p + 8];
        if (p + 8 + size > block.size()) break;
        switch (type) {
            case 0x0001:
                parse_dataspace(body, size, obj.dataset.dims);
                break;
            case 0x0003:
                obj.dataset.type = parse_datatype(body, size);
                break;
            case 0x0008: {
                Dataset& d = obj.dataset;
                const int version = body[0];
                obj.is_dataset = true;
                if (version == 3) {
                    d.layout = body[1];
                    if (d.layout == 0) {
                        const uint16_t n = get<uint16_t>(body + 2);
                        d.compact.assign(body + 4, body + 4 + n);
                    } else if (d.layout == 1) {
                        d.address = get<uint64_t>(body + 2);
                        d.storage_size = get<uint64_t>(body + 10);
                    } else if (d.layout == 2) {
                        const int rank = body[2];
                        d.address = get<uint64_t>(body + 3);
                        for (int i = 0; i + 1 < rank; ++i) d.chunk_dims.push_back(get<uint32_t>(body + 11 + 4 * i));
                    } else throw Error("unsupported data layout class");
                } else if (version == 1 || version == 2) {
                    const int rank = body[1];
                    d.layout = body[2];
                    size_t off = 8;
                    if (d.layout != 0) {
                        d.address = get<uint64_t>(body + off);
                        off += 8;
                    }
                    std::vector<uint32_t> dims;
                    for (int i = 0; i < rank; ++i) dims.push_back(get<uint32_t>(body + off + 4 * i));
                    off += 4 * (size_t)rank;
                    if (d.layout == 0) {
                        const uint32_t n = get<uint32_t>(body + off);
                        d.compact.assign(body + off + 4, body + off + 4 + n);
                    } else if (d.layout == 2) {
                        for (int i = 0; i + 1 < rank; ++i) d.chunk_dims.push_back(dims[i]);
                    }
                } else if (version == 4) {
                    // layout version 4 (libver "latest"): compact and contiguous as in version 3; chunked: flags,
                    // dimensionality, width of a dimension size, the sizes (the last one is the element size), the kind of
                    // chunk index with its parameters, the index's address
                    d.layout = body[1];
                    if (d.layout == 0) {
                        const uint16_t n = get<uint16_t>(body + 2);
                        d.compact.assign(body + 4, body + 4 + n);
                    } else if (d.layout == 1) {
                        d.address = get<uint64_t>(body + 2);
                        d.storage_size = get<uint64_t>(body + 10);
                    } else if (d.layout == 2) {
                        const uint8_t cflags = body[2];
                        const int rank = body[3], enc = body[4];
                        size_t off = 5;
                        for (int i = 0; i < rank; ++i, off += enc) {
                            uint64_t v = 0;
                            std::memcpy(&v, body + off, enc);
                            if (i + 1 < rank) d.chunk_dims.push_back(v);
                        }
                        d.chunk_index = body[off++];
                        if (d.chunk_index == 1) {  // a single chunk
                            if (cflags & 2) {
                                d.single_chunk_size = get<uint64_t>(body + off);
                                d.single_filter_mask = get<uint32_t>(body + off + 8);
                                off += 12;
                            }
                        } else if (d.chunk_index == 2) {  // implicit: all chunks allocated, in order, no filters
                        } else if (d.chunk_index == 3) {  // fixed array
                            off += 1;                      // page bits
                        } else {
                            throw Error(filename + ": chunk index type " + std::to_string(d.chunk_index) +
                                        " (extensible array / version-2 B-tree) is not supported by this reader");
                        }
                        d.address = get<uint64_t>(body + off);
                    } else throw Error("unsupported data layout class");
                } else throw Error("data layout message version " + std::to_string(version) + " is not supported");
                break;
            }
            case 0x000B: {
                const int version = body[0], nf = body[1];
                size_t off = (version == 1) ? 8 : 2;
                for (int i = 0; i < nf; ++i) {
                    const uint16_t id = get<uint16_t>(body + off);
                    uint16_t name_len = 0;
                    size_t q = off + 2;
                    if (version == 1 || id >= 256) {
                        name_len = get<uint16_t>(body + q);
                        q += 2;
                    }
                    q += 2;  // flags
                    const uint16_t ncd = get<uint16_t>(body + q);
                    q += 2;
                    q += (version == 1) ? pad8(name_len) : name_len;
                    q += 4 * (size_t)ncd;
                    if (version == 1 && (ncd & 1)) q += 4;
                    obj.dataset.filters.push_back(id);
                    off = q;
                }
                break;
            }
            case 0x000C: {
                std::string name;
                Attribute a = parse_attribute(body, size, name);
                obj.attributes[name] = std::move(a);
                break;
            }
            case 0x0010:
                more.emplace_back(get<uint64_t>(body), get<uint64_t>(body + 8));
                break;
            case 0x0011: {
                obj.is_group = true;
                const uint64_t btree = get<uint64_t>(body), heap = get<uint64_t>(body + 8);
                uint8_t h[32];
                read_at(heap, h, 32);
                if (std::memcmp(h, "HEAP", 4) != 0) throw Error("bad local heap in " + filename);
                walk_group_btree(btree, get<uint64_t>(h + 24), obj.links);
                break;
            }
            case 0x0002: {
                // link info (new-style group): version, flags, [max creation index], fractal heap address, name index address.
                // The links of a small group are link messages in the header ("compact"); a defined heap address means
                // "dense" storage (fractal heap + version-2 B-tree), which this reader does not follow.
                obj.is_group = true;
                const size_t off = 2 + ((body[1] & 1) ? 8 : 0);
                if (size >= off + 8 && get<uint64_t>(body + off) != kUndefined)
                    throw Error(filename + " keeps a group's links in a fractal heap (more than 8 links with libver 'latest'), which this reader does not support");
                break;
            }
            case 0x0006: {
                // link message: version 1, flags (bits 0-1: width of the name length, 2: creation order, 3: link type,
                // 4: character set), [type], [creation order], [charset], name length, name; hard link: object header address
                obj.is_group = true;
                const uint8_t lf = body[1];
                size_t off = 2;
                uint8_t link_type = 0;
                if (lf & 0x08) link_type = body[off++];
                if (lf & 0x04) off += 8;
                if (lf & 0x10) off += 1;
                const int w = 1 << (lf & 3);
                uint64_t name_len = 0;
                std::memcpy(&name_len, body + off, w);
                off += w;
                if (off + name_len + 8 > size) break;
                const std::string name(reinterpret_cast<const char*>(body + off), name_len);
                off += name_len;
                if (link_type == 0) obj.links[name] = get<uint64_t>(body + off);  // soft / external links are skipped
                break;
            }
            default:
                break;
        }
        p += 8 + pad8(size);
    }
}

Object File::parse_object(uint64_t header_address) const {
    uint8_t head[16];
    read_at(header_address, head, 16);
    if (std::memcmp(head, "OHDR", 4) == 0) return parse_object_v2(header_address);
    if (head[0] != 1) throw Error("bad object header in " + filename);
    const uint32_t size = get<uint32_t>(head + 8);
    Object obj;
    std::vector<std::pair<uint64_t, uint64_t>> blocks = {{header_address + 16, size}};
    for (size_t b = 0; b < blocks.size(); ++b) {
        const std::vector<uint8_t> block = bytes_at(blocks[b].first, blocks[b].second);
        parse_messages(block, obj, blocks);
    }
    return obj;
}

// Version-2 object header (libver "latest"): "OHDR", version 2, flags, [four 4-byte times if flags & 0x20], [max compact /
// min dense attribute counts if flags & 0x10], size of chunk 0 in 1 << (flags & 3) bytes, then messages
// {type u8, size u16, flags u8, [creation order u16 if flags & 0x04]} ... and a 4-byte checksum; continuation blocks are
// "OCHK" + messages + checksum. Messages are not padded. The bodies are those of version-1 headers, plus the link message
// that replaces symbol tables; they are repacked as version-1 records so that one message parser serves both.
Object File::parse_object_v2(uint64_t header_address) const {
    uint8_t head[6];
    read_at(header_address, head, 6);
    if (head[4] != 2) throw Error("bad version-2 object header in " + filename);
    const uint8_t flags = head[5];
    uint64_t p = header_address + 6;
    if (flags & 0x20) p += 16;
    if (flags & 0x10) p += 4;
    const int size_bytes = 1 << (flags & 3);
    uint8_t sz[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    read_at(p, sz, size_bytes);
    const uint64_t chunk0 = get<uint64_t>(sz);
    p += size_bytes;
    const size_t prefix = (flags & 0x04) ? 6 : 4;
    Object obj;
    std::vector<std::pair<uint64_t, uint64_t>> blocks = {{p, chunk0}};
    for (size_t b = 0; b < blocks.size(); ++b) {
        const std::vector<uint8_t> block = bytes_at(blocks[b].first, blocks[b].second);
        std::vector<uint8_t> v1;
        size_t q = 0;
        while (q + prefix <= block.size()) {
            const uint8_t type = block[q];
            const uint16_t size = get<uint16_t>(&block[q + 1]);
            const uint8_t mflags = block[q + 3];
            if (q + prefix + size > block.size()) break;
            const uint8_t* body = &block[q + prefix];
            if (type == 0x10) {
                // continuation: the block starts with "OCHK" and ends with a checksum
                blocks.emplace_back(get<uint64_t>(body) + 4, get<uint64_t>(body + 8) - 8);
            } else if (type != 0) {
                const uint16_t padded = (uint16_t)pad8(size);
                const uint8_t rec[8] = {type, 0, (uint8_t)(padded & 0xff), (uint8_t)(padded >> 8), mflags, 0, 0, 0};
                v1.insert(v1.end(), rec, rec + 8);
                v1.insert(v1.end(), body, body + size);
                v1.resize(v1.size() + (padded - size), 0);
            }
            q += prefix + size;
        }
        std::vector<std::pair<uint64_t, uint64_t>> none;
        parse_messages(v1, obj, none);
    }
    return obj;
}

uint64_t File::resolve(const std::string& path) const {
    uint64_t addr = root_header;
    size_t p = 0;
    while (p < path.size()) {
        while (p < path.size() && path[p] == '/') ++p;
        if (p >= path.size()) break;
        const size_t q = path.find('/', p);
        const std::string part = path.substr(p, q == std::string::npos ? std::string::npos : q - p);
        const Object o = parse_object(addr);
        auto it = o.links.find(part);
        if (it == o.links.end()) return kUndefined;
        addr = it->second;
        p = (q == std::string::npos) ? path.size() : q + 1;
    }
    return addr;
}

bool File::exists(const std::string& path) const { return resolve(path) != kUndefined; }

Object File::object(const std::string& path) const {
    const uint64_t addr = resolve(path);
    if (addr == kUndefined) throw Error("no object '" + path + "' in " + filename);
    return parse_object(addr);
}

std::vector<std::string> File::list(const std::string& group_path) const {
    std::vector<std::string> names;
    for (const auto& kv : object(group_path).links) names.push_back(kv.first);
    return names;
}

bool File::has_attribute(const std::string& object_path, const std::string& name) const {
    const Object o = object(object_path);
    return o.attributes.count(name) != 0;
}

template <typename T>
T File::attribute(const std::string& object_path, const std::string& name) const {
    const Object o = object(object_path);
    auto it = o.attributes.find(name);
    if (it == o.attributes.end()) throw Error("no attribute '" + name + "' on '" + object_path + "' in " + filename);
    if (it->second.raw.size() < it->second.type.size) throw Error("empty attribute '" + name + "'");
    return convert_one<T>(it->second.raw.data(), it->second.type);
}

std::string File::string_attribute(const std::string& object_path, const std::string& name) const {
    const Object o = object(object_path);
    auto it = o.attributes.find(name);
    if (it == o.attributes.end()) throw Error("no attribute '" + name + "' on '" + object_path + "' in " + filename);
    if (it->second.strings.empty()) throw Error("attribute '" + name + "' is not a string");
    return it->second.strings[0];
}

void File::read_chunked(const Dataset& d, std::vector<uint8_t>& whole) const {
    const size_t rank = d.dims.size();
    if (d.chunk_dims.size() != rank) throw Error("chunk rank mismatch in " + filename);
    const size_t elem = d.type.size;
    whole.assign(d.elements() * elem, 0);
    uint64_t chunk_elems = 1;
    for (uint64_t c : d.chunk_dims) chunk_elems *= c;

    // one chunk: read, undo the filters, scatter its rows into the row-major whole (last dimension contiguous)
    auto place = [&](const std::vector<uint64_t>& off, uint64_t address, uint64_t stored, uint32_t mask) {
        std::vector<uint8_t> chunk = bytes_at(address, stored);
        for (size_t f = d.filters.size(); f-- > 0;) {  // undo the pipeline in reverse order
            if (mask & (1u << f)) continue;
            if (d.filters[f] == 1) {
                std::vector<uint8_t> out(chunk_elems * elem);
                uLongf out_len = out.size();
                if (uncompress(out.data(), &out_len, chunk.data(), chunk.size()) != Z_OK)
                    throw Error("cannot inflate a chunk of " + filename);
                out.resize(out_len);
                chunk.swap(out);
            } else if (d.filters[f] == 2) {
                unshuffle(chunk, elem);
            } else {
                throw Error("unsupported filter " + std::to_string(d.filters[f]) + " in " + filename);
            }
        }
        if (chunk.size() < chunk_elems * elem) throw Error("short chunk in " + filename);
        std::vector<uint64_t> idx(rank, 0);
        const uint64_t last = d.chunk_dims[rank - 1];
        const uint64_t rows = chunk_elems / last;
        for (uint64_t r = 0; r < rows; ++r) {
            uint64_t rem = r;
            for (size_t k = rank - 1; k-- > 0;) {
                idx[k] = rem % d.chunk_dims[k];
                rem /= d.chunk_dims[k];
            }
            bool inside = true;
            uint64_t flat = 0;
            for (size_t k = 0; k + 1 < rank; ++k) {
                const uint64_t g = off[k] + idx[k];
                if (g >= d.dims[k]) inside = false;
                flat = flat * d.dims[k] + g;
            }
            if (!inside || off[rank - 1] >= d.dims[rank - 1]) continue;
            const uint64_t n = std::min<uint64_t>(last, d.dims[rank - 1] - off[rank - 1]);
            std::memcpy(&whole[(flat * d.dims[rank - 1] + off[rank - 1]) * elem], &chunk[r * last * elem], n * elem);
        }
    };
    if (d.address == kUndefined) return;  // nothing allocated: fill value

    if (d.chunk_index != 0) {
        // layout version 4. Chunks are numbered in row-major order of the chunk grid.
        std::vector<uint64_t> grid(rank);
        uint64_t nchunks = 1;
        for (size_t k = 0; k < rank; ++k) {
            grid[k] = (d.dims[k] + d.chunk_dims[k] - 1) / d.chunk_dims[k];
            nchunks *= grid[k];
        }
        auto offsets_of = [&](uint64_t c) {
            std::vector<uint64_t> off(rank);
            for (size_t k = rank; k-- > 0;) {
                off[k] = (c % grid[k]) * d.chunk_dims[k];
                c /= grid[k];
            }
            return off;
        };
        const uint64_t plain = chunk_elems * elem;
        if (d.chunk_index == 1) {  // one chunk holds the whole dataset
            place(std::vector<uint64_t>(rank, 0), d.address, d.single_chunk_size ? d.single_chunk_size : plain,
                  d.single_chunk_size ? d.single_filter_mask : 0);
        } else if (d.chunk_index == 2) {  // implicit: every chunk allocated, one after the other, unfiltered
            for (uint64_t c = 0; c < nchunks; ++c) place(offsets_of(c), d.address + c * plain, plain, 0);
        } else {  // fixed array: header "FAHD" -> data block "FADB" -> one element per chunk
            uint8_t h[28];
            read_at(d.address, h, 28);
            if (std::memcmp(h, "FAHD", 4) != 0) throw Error("bad fixed-array header in " + filename);
            const int client = h[5], entry = h[6], page_bits = h[7];
            const uint64_t nelem = get<uint64_t>(h + 8), block = get<uint64_t>(h + 16);
            if (block == kUndefined) return;
            if (nelem > (uint64_t(1) << page_bits))
                throw Error(filename + ": paged fixed-array chunk index (more than " + std::to_string(1u << page_bits) + " chunks) is not supported by this reader");
            const std::vector<uint8_t> elems = bytes_at(block + 14, (size_t)nelem * entry);  // "FADB", version, client, header address
            for (uint64_t c = 0; c < std::min(nelem, nchunks); ++c) {
                const uint8_t* e = &elems[(size_t)c * entry];
                const uint64_t address = get<uint64_t>(e);
                if (address == kUndefined) continue;
                if (client == 0) {
                    place(offsets_of(c), address, plain, 0);
                } else {  // filtered chunks: address, stored size (entry - 12 bytes wide), filter mask
                    uint64_t stored = 0;
                    std::memcpy(&stored, e + 8, entry - 12);
                    place(offsets_of(c), address, stored, get<uint32_t>(e + entry - 4));
                }
            }
        }
        return;
    }

    // depth-first over the v1 chunk B-tree (node type 1); key = chunk size (4), filter mask (4), offsets (rank + 1) x 8
    std::vector<uint64_t> stack = {d.address};
    const size_t key_bytes = 8 + 8 * (rank + 1);
    while (!stack.empty()) {
        const uint64_t node = stack.back();
        stack.pop_back();
        if (node == kUndefined) continue;
        uint8_t head[24];
        read_at(node, head, 24);
        if (std::memcmp(head, "TREE", 4) != 0 || head[4] != 1) throw Error("bad chunk B-tree node in " + filename);
        const int level = head[5], entries = get<uint16_t>(head + 6);
        const std::vector<uint8_t> body = bytes_at(node + 24, (size_t)entries * (key_bytes + 8) + key_bytes);
        for (int i = 0; i < entries; ++i) {
            const uint8_t* key = &body[(size_t)i * (key_bytes + 8)];
            const uint64_t child = get<uint64_t>(key + key_bytes);
            if (level > 0) {
                stack.push_back(child);
                continue;
            }
            std::vector<uint64_t> off(rank);
            for (size_t r = 0; r < rank; ++r) off[r] = get<uint64_t>(key + 8 + 8 * r);
            place(off, child, get<uint32_t>(key), get<uint32_t>(key + 4));
        }
    }
}

void File::read_raw(const Dataset& d, uint64_t first_element, uint64_t count, std::vector<uint8_t>& out) const {
    const size_t elem = d.type.size;
    out.resize(count * elem);
    if (count == 0) return;
    if (d.layout == 1) {
        if (d.address == kUndefined) {
            std::fill(out.begin(), out.end(), 0);  // never written: fill value (0)
            return;
        }
        read_at(d.address + first_element * elem, out.data(), out.size());
    } else if (d.layout == 0) {
        if ((first_element + count) * elem > d.compact.size()) throw Error("compact dataset too small in " + filename);
        std::memcpy(out.data(), &d.compact[first_element * elem], out.size());
    } else if (d.layout == 2) {
        std::vector<uint8_t> whole;
        read_chunked(d, whole);
        std::memcpy(out.data(), &whole[first_element * elem], out.size());
    } else {
        throw Error("object is not a dataset in " + filename);
    }
}

template <typename T>
std::vector<T> File::read_rows(const std::string& path, uint64_t row0, uint64_t nrows) const {
    const Object o = object(path);
    if (!o.is_dataset) throw Error("'" + path + "' is not a dataset in " + filename);
    const Dataset& d = o.dataset;
    uint64_t row_elems = 1;
    for (size_t k = 1; k < d.dims.size(); ++k) row_elems *= d.dims[k];
    const uint64_t total_rows = d.dims.empty() ? 1 : d.dims[0];
    if (row0 + nrows > total_rows) throw Error("rows out of range for '" + path + "' in " + filename);
    std::vector<uint8_t> raw;
    read_raw(d, row0 * row_elems, nrows * row_elems, raw);
    std::vector<T> out(nrows * row_elems);
    for (uint64_t i = 0; i < out.size(); ++i) out[i] = convert_one<T>(&raw[i * d.type.size], d.type);
    return out;
}

template <typename T>
std::vector<T> File::read(const std::string& path) const {
    const Object o = object(path);
    const uint64_t rows = o.dataset.dims.empty() ? 1 : o.dataset.dims[0];
    return read_rows<T>(path, 0, rows);
}

uint64_t File::contiguous_offset(const std::string& path) const {
    const Object o = object(path);
    if (!o.is_dataset || o.dataset.layout != 1 || !o.dataset.filters.empty() || o.dataset.address == kUndefined) return kUndefined;
    return base + o.dataset.address;
}

template std::vector<double> File::read<double>(const std::string&) const;
template std::vector<float> File::read<float>(const std::string&) const;
template std::vector<uint64_t> File::read<uint64_t>(const std::string&) const;
template std::vector<int32_t> File::read<int32_t>(const std::string&) const;
template std::vector<int64_t> File::read<int64_t>(const std::string&) const;
template std::vector<double> File::read_rows<double>(const std::string&, uint64_t, uint64_t) const;
template std::vector<float> File::read_rows<float>(const std::string&, uint64_t, uint64_t) const;
template std::vector<int32_t> File::read_rows<int32_t>(const std::string&, uint64_t, uint64_t) const;
template double File::attribute<double>(const std::string&, const std::string&) const;
template uint64_t File::attribute<uint64_t>(const std::string&, const std::string&) const;
template int32_t File::attribute<int32_t>(const std::string&, const std::string&) const;
template int64_t File::attribute<int64_t>(const std::string&, const std::string&) const;

// ---- writer -------------------------------------------------------------------------------------------------------------

namespace {

template <typename T>
void put(std::vector<uint8_t>& v, T x) {
    const uint8_t* p = reinterpret_cast<const uint8_t*>(&x);
    v.insert(v.end(), p, p + sizeof(T));
}
void put_bytes(std::vector<uint8_t>& v, const void* p, size_t n) {
    const uint8_t* b = static_cast<const uint8_t*>(p);
    v.insert(v.end(), b, b + n);
}
void pad_to8(std::vector<uint8_t>& v) {
    while (v.size() % 8) v.push_back(0);
}

// datatype messages (version 1), little-endian
std::vector<uint8_t> type_integer(uint32_t bytes, bool is_signed) {
    std::vector<uint8_t> m = {0x10, (uint8_t)(is_signed ? 0x08 : 0x00), 0, 0};
    put<uint32_t>(m, bytes);
    put<uint16_t>(m, 0);
    put<uint16_t>(m, (uint16_t)(8 * bytes));
    return m;
}
std::vector<uint8_t> type_float(uint32_t bytes) {
    // class bits: byte order LE, padding 0, mantissa normalisation 2 (implied msb), sign location
    std::vector<uint8_t> m = {0x11, 0x20, (uint8_t)(bytes == 4 ? 31 : 63), 0};
    put<uint32_t>(m, bytes);
    put<uint16_t>(m, 0);
    put<uint16_t>(m, (uint16_t)(8 * bytes));
    if (bytes == 4) {
        m.insert(m.end(), {23, 8, 0, 23});
        put<uint32_t>(m, 127);
    } else {
        m.insert(m.end(), {52, 11, 0, 52});
        put<uint32_t>(m, 1023);
    }
    return m;
}
std::vector<uint8_t> type_string(uint32_t bytes) {
    std::vector<uint8_t> m = {0x13, 0x00, 0, 0};  // null-terminated, ASCII
    put<uint32_t>(m, bytes);
    return m;
}
std::vector<uint8_t> space_simple(const std::vector<uint64_t>& dims) {
    std::vector<uint8_t> m = {1, (uint8_t)dims.size(), 0, 0, 0, 0, 0, 0};
    for (uint64_t d : dims) put<uint64_t>(m, d);
    return m;
}

void append_message(std::vector<uint8_t>& hdr, uint16_t type, const std::vector<uint8_t>& body, uint8_t flags = 0) {
    put<uint16_t>(hdr, type);
    put<uint16_t>(hdr, (uint16_t)pad8(body.size()));
    hdr.push_back(flags);
    hdr.insert(hdr.end(), {0, 0, 0});
    put_bytes(hdr, body.data(), body.size());
    pad_to8(hdr);
}

}  // namespace

Writer::Node::~Node() {
    for (auto& kv : children) delete kv.second;
}

Writer::Writer(const std::string& path) : filename(path) {}

Writer::~Writer() {
    try {
        if (!closed) close();
        flush();
    } catch (...) {
    }
    if (append_fp) std::fclose(append_fp);
}

Writer::Node* Writer::find(const std::string& path, bool parent_only, std::string* leaf) {
    Node* n = &root;
    size_t p = 0;
    std::vector<std::string> parts;
    while (p < path.size()) {
        while (p < path.size() && path[p] == '/') ++p;
        if (p >= path.size()) break;
        const size_t q = path.find('/', p);
        parts.push_back(path.substr(p, q == std::string::npos ? std::string::npos : q - p));
        p = (q == std::string::npos) ? path.size() : q + 1;
    }
    const size_t upto = parent_only ? (parts.empty() ? 0 : parts.size() - 1) : parts.size();
    for (size_t i = 0; i < upto; ++i) {
        auto it = n->children.find(parts[i]);
        if (it == n->children.end()) throw Error("no such object: " + path);
        n = it->second;
    }
    if (leaf) *leaf = parts.empty() ? "" : parts.back();
    return n;
}

void Writer::create_group(const std::string& path) {
    std::string leaf;
    Node* parent = find(path, true, &leaf);
    if (leaf.empty() || parent->children.count(leaf)) throw Error("cannot create group " + path);
    parent->children[leaf] = new Node();
}

void Writer::add_dataset(const std::string& path, const std::vector<uint64_t>& dims, const std::vector<uint8_t>& datatype,
                         const void* data, uint64_t bytes) {
    std::string leaf;
    Node* parent = find(path, true, &leaf);
    if (leaf.empty() || !parent->is_group || parent->children.count(leaf)) throw Error("cannot create dataset " + path);
    Node* n = new Node();
    n->is_group = false;
    n->datatype = datatype;
    n->dataspace = space_simple(dims);
    n->owned.assign(static_cast<const uint8_t*>(data), static_cast<const uint8_t*>(data) + bytes);
    n->data_bytes = bytes;
    parent->children[leaf] = n;
}

static uint64_t count_of(const std::vector<uint64_t>& dims) {
    uint64_t n = 1;
    for (uint64_t d : dims) n *= d;
    return n;
}

void Writer::write_dataset(const std::string& path, const std::vector<uint64_t>& dims, const double* data) {
    add_dataset(path, dims, type_float(8), data, count_of(dims) * 8);
}
void Writer::write_dataset(const std::string& path, const std::vector<uint64_t>& dims, const float* data) {
    add_dataset(path, dims, type_float(4), data, count_of(dims) * 4);
}
void Writer::write_dataset(const std::string& path, const std::vector<uint64_t>& dims, const int32_t* data) {
    add_dataset(path, dims, type_integer(4, true), data, count_of(dims) * 4);
}
void Writer::write_dataset(const std::string& path, const std::vector<uint64_t>& dims, const uint64_t* data) {
    add_dataset(path, dims, type_integer(8, false), data, count_of(dims) * 8);
}

void Writer::create_extendible(const std::string& path, bool is_double, const std::vector<uint64_t>& row_shape, uint64_t chunk_rows) {
    if (closed) throw Error("extendible datasets must be created before the file is laid out");
    if (chunk_rows == 0) throw Error("chunk_rows must be positive");
    std::string leaf;
    Node* parent = find(path, true, &leaf);
    if (leaf.empty() || !parent->is_group || parent->children.count(leaf)) throw Error("cannot create dataset " + path);
    Node* n = new Node();
    n->is_group = false;
    n->extendible = true;
    n->elem_size = is_double ? 8 : 4;
    n->datatype = is_double ? type_float(8) : type_integer(4, true);
    n->row_shape = row_shape;
    n->chunk_rows = chunk_rows;
    n->row_bytes = n->elem_size * count_of(row_shape);
    const uint64_t chunk_bytes = n->row_bytes * chunk_rows;
    if (chunk_bytes == 0 || chunk_bytes > 0xffffffffULL) throw Error("chunk of " + path + " is empty or larger than 4 GiB");
    // dataspace message version 1 with maximum dimensions: the first one unlimited
    std::vector<uint8_t> sp = {1, (uint8_t)(row_shape.size() + 1), 1, 0, 0, 0, 0, 0};
    put<uint64_t>(sp, 0);
    for (uint64_t d : row_shape) put<uint64_t>(sp, d);
    put<uint64_t>(sp, kUndefined);
    for (uint64_t d : row_shape) put<uint64_t>(sp, d);
    n->dataspace = sp;
    parent->children[leaf] = n;
    extendibles.push_back(n);
}

void Writer::write_at(uint64_t address, const void* data, size_t bytes) {
    if (!append_fp) {
        append_fp = std::fopen(filename.c_str(), "r+b");
        if (!append_fp) throw Error("cannot reopen " + filename + " for appending");
    }
    if (fseeko(append_fp, (off_t)address, SEEK_SET) != 0 || std::fwrite(data, 1, bytes, append_fp) != bytes)
        throw Error("short write on " + filename);
}

void Writer::append_rows(const std::string& path, const void* rows, uint64_t nrows) {
    Node* n = find(path, false, nullptr);
    if (!n->extendible) throw Error(path + " is not an extendible dataset");
    const uint8_t* src = static_cast<const uint8_t*>(rows);
    if (!closed) {
        n->pending.insert(n->pending.end(), src, src + nrows * n->row_bytes);
        return;
    }
    append_to(n, src, nrows);
}

void Writer::append_to(Node* n, const uint8_t* src, uint64_t nrows) {
    const uint64_t chunk_bytes = n->row_bytes * n->chunk_rows;
    while (nrows) {
        const uint64_t chunk = n->rows / n->chunk_rows, within = n->rows % n->chunk_rows;
        if (chunk == n->chunk_address.size()) {  // a new chunk at the end of the file, fill value (0) behind the rows
            const std::vector<uint8_t> zeros(chunk_bytes, 0);
            n->chunk_address.push_back(eof);
            write_at(eof, zeros.data(), zeros.size());
            eof += chunk_bytes;
        }
        const uint64_t take = std::min(nrows, n->chunk_rows - within);
        write_at(n->chunk_address[chunk] + within * n->row_bytes, src, take * n->row_bytes);
        src += take * n->row_bytes;
        n->rows += take;
        nrows -= take;
        n->dirty = true;
    }
}

// Version-1 B-tree of node type 1 over the dataset's chunks, written at the end of the file; returns the root's address.
// Key: chunk size in bytes, filter mask, the chunk's offset in every dimension plus one zero for the element; a node
// holds up to 64 children (the default "indexed storage" rank of superblock version 0) and is allocated at full size,
// as the library reads it; the last key of a node bounds its last child from above.
uint64_t Writer::write_chunk_index(Node* n) {
    constexpr size_t kFan = 64;
    const size_t rank = n->row_shape.size() + 1;
    const size_t key_bytes = 8 + 8 * (rank + 1);
    const size_t node_bytes = 24 + kFan * (key_bytes + 8) + key_bytes;
    const uint32_t chunk_bytes = (uint32_t)(n->row_bytes * n->chunk_rows);
    auto key_of = [&](uint64_t first_row, std::vector<uint8_t>& out) {
        put<uint32_t>(out, chunk_bytes);
        put<uint32_t>(out, 0);
        put<uint64_t>(out, first_row);
        for (size_t k = 1; k <= rank; ++k) put<uint64_t>(out, 0);  // other dimensions start at 0; the element entry is 0
    };
    struct Entry {
        uint64_t first_row, address;
    };
    std::vector<Entry> level_entries;
    for (size_t c = 0; c < n->chunk_address.size(); ++c) level_entries.push_back({c * n->chunk_rows, n->chunk_address[c]});
    const uint64_t end_row = n->chunk_address.size() * n->chunk_rows;
    int level = 0;
    while (true) {
        const size_t nnodes = std::max<size_t>(1, (level_entries.size() + kFan - 1) / kFan);
        std::vector<Entry> parents;
        const uint64_t first_node = eof;
        for (size_t b = 0; b < nnodes; ++b) {
            const size_t lo = b * kFan, hi = std::min(level_entries.size(), lo + kFan);
            std::vector<uint8_t> node;
            put_bytes(node, "TREE", 4);
            node.push_back(1);
            node.push_back((uint8_t)level);
            put<uint16_t>(node, (uint16_t)(hi - lo));
            put<uint64_t>(node, b ? first_node + (b - 1) * node_bytes : kUndefined);              // left sibling
            put<uint64_t>(node, b + 1 < nnodes ? first_node + (b + 1) * node_bytes : kUndefined);  // right sibling
            for (size_t i = lo; i < hi; ++i) {
                key_of(level_entries[i].first_row, node);
                put<uint64_t>(node, level_entries[i].address);
            }
            key_of(hi < level_entries.size() ? level_entries[hi].first_row : end_row, node);
            node.resize(node_bytes, 0);
            write_at(eof, node.data(), node.size());
            parents.push_back({lo < level_entries.size() ? level_entries[lo].first_row : 0, eof});
            eof += node_bytes;
        }
        if (nnodes == 1) return first_node;
        level_entries.swap(parents);
        ++level;
    }
}

void Writer::flush() {
    if (!closed) return;
    bool any = false;
    for (Node* n : extendibles) {
        if (!n->dirty) continue;
        const uint64_t root_node = write_chunk_index(n);
        write_at(n->index_field, &root_node, 8);
        write_at(n->extent_field, &n->rows, 8);
        n->dirty = false;
        any = true;
    }
    if (any) {
        write_at(40, &eof, 8);  // end-of-file address of the superblock
        std::fflush(append_fp);
    }
}

void Writer::add_attribute(const std::string& object_path, const std::string& name, const std::vector<uint8_t>& datatype,
                           const void* data, uint32_t bytes) {
    Node* n = find(object_path, false, nullptr);
    // attribute message version 1: name, datatype and dataspace each padded to 8 bytes
    std::vector<uint8_t> m = {1, 0};
    const std::vector<uint8_t> space = {1, 0, 0, 0, 0, 0, 0, 0};  // scalar
    put<uint16_t>(m, (uint16_t)(name.size() + 1));
    put<uint16_t>(m, (uint16_t)datatype.size());
    put<uint16_t>(m, (uint16_t)space.size());
    put_bytes(m, name.c_str(), name.size() + 1);
    pad_to8(m);
    put_bytes(m, datatype.data(), datatype.size());
    pad_to8(m);
    put_bytes(m, space.data(), space.size());
    pad_to8(m);
    put_bytes(m, data, bytes);
    n->attribute_messages.emplace_back(name, m);
}

void Writer::set_attribute(const std::string& p, const std::string& name, double v) { add_attribute(p, name, type_float(8), &v, 8); }
void Writer::set_attribute(const std::string& p, const std::string& name, uint64_t v) { add_attribute(p, name, type_integer(8, false), &v, 8); }
void Writer::set_attribute(const std::string& p, const std::string& name, int32_t v) { add_attribute(p, name, type_integer(4, true), &v, 4); }
void Writer::set_attribute(const std::string& p, const std::string& name, const std::string& v) {
    add_attribute(p, name, type_string((uint32_t)v.size() + 1), v.c_str(), (uint32_t)v.size() + 1);
}

void Writer::close() {
    if (closed) return;
    closed = true;
    constexpr int kLeafK = 32;  // up to 64 links per symbol node (written into the superblock)
    constexpr int kInternalK = 16;

    // pass 1: sizes and addresses, depth first; the layout of every object is known before anything is written
    uint64_t cursor = 96;  // superblock (24 + 32 fixed/addresses + 40 root entry)
    std::vector<Node*> order;
    std::vector<std::vector<uint8_t>> headers;
    struct Plan {
        Node* n;
        uint64_t header_size;
    };
    std::vector<Node*> stack = {&root};
    while (!stack.empty()) {
        Node* n = stack.back();
        stack.pop_back();
        order.push_back(n);
        for (auto it = n->children.rbegin(); it != n->children.rend(); ++it) stack.push_back(it->second);
    }
    auto attr_bytes = [](const Node* n) {
        uint64_t b = 0;
        for (const auto& a : n->attribute_messages) b += 8 + pad8(a.second.size());
        return b;
    };
    for (Node* n : order) {
        if (n->children.size() > 2 * kLeafK) throw Error("too many links in one group for this writer");
        n->header_address = cursor;
        if (n->is_group) {
            cursor += 16 + (8 + 16) + attr_bytes(n);  // prefix + symbol table message + attributes
            n->btree = cursor;
            cursor += 24 + (2 * kInternalK + 1) * 8 + 2 * kInternalK * 8;  // full-size node as libhdf5 allocates it
            uint64_t names = 8;  // offset 0 holds the empty string
            for (const auto& kv : n->children) names += pad8(kv.first.size() + 1);
            n->heap = cursor;
            cursor += 32;
            n->heap_data = cursor;
            n->data_bytes = pad8(names + 16);  // + a free block
            cursor += n->data_bytes;
            n->snod = cursor;
            cursor += 8 + 2 * kLeafK * 40;
        } else if (n->extendible) {
            // dataspace, datatype, fill value (version 2 with the value), chunked layout (version 3)
            const uint64_t fill = 8 + pad8(8 + n->elem_size), layout = 8 + pad8(3 + 8 + 4 * (n->row_shape.size() + 2));
            cursor += 16 + (8 + pad8(n->dataspace.size())) + (8 + pad8(n->datatype.size())) + fill + layout + attr_bytes(n);
            cursor = pad8(cursor);
        } else {
            cursor += 16 + (8 + pad8(n->dataspace.size())) + (8 + pad8(n->datatype.size())) + (8 + 8) /*fill*/ + (8 + 24) /*layout*/ + attr_bytes(n);
            cursor = pad8(cursor);
            n->data_address = cursor;
            cursor += pad8(n->owned.size());
        }
    }
    eof = cursor;

    std::vector<uint8_t> f(eof, 0);
    auto at = [&](uint64_t a) { return f.data() + a; };
    auto store = [&](uint64_t a, const std::vector<uint8_t>& v) { std::memcpy(at(a), v.data(), v.size()); };

    // superblock version 0
    {
        std::vector<uint8_t> sb;
        put_bytes(sb, kSignature, 8);
        sb.insert(sb.end(), {0, 0, 0, 0, 0, 8, 8, 0});
        put<uint16_t>(sb, kLeafK);
        put<uint16_t>(sb, kInternalK);
        put<uint32_t>(sb, 0);
        put<uint64_t>(sb, 0);           // base address
        put<uint64_t>(sb, kUndefined);  // free-space info
        put<uint64_t>(sb, eof);
        put<uint64_t>(sb, kUndefined);  // driver info
        put<uint64_t>(sb, 0);           // root: link name offset
        put<uint64_t>(sb, root.header_address);
        put<uint32_t>(sb, 1);           // cache type 1: group, scratch = B-tree + heap addresses
        put<uint32_t>(sb, 0);
        put<uint64_t>(sb, root.btree);
        put<uint64_t>(sb, root.heap);
        store(0, sb);
    }

    for (Node* n : order) {
        std::vector<uint8_t> msgs;
        if (n->is_group) {
            std::vector<uint8_t> st;
            put<uint64_t>(st, n->btree);
            put<uint64_t>(st, n->heap);
            append_message(msgs, 0x0011, st);
        } else if (n->extendible) {
            n->extent_field = n->header_address + 16 + msgs.size() + 8 + 8;  // dims[0] inside the dataspace message
            append_message(msgs, 0x0001, n->dataspace);
            append_message(msgs, 0x0003, n->datatype, 1);
            std::vector<uint8_t> fill = {2, 3, 2, 1};  // version 2: allocate incrementally, write if set, value defined (0)
            put<uint32_t>(fill, n->elem_size);
            fill.resize(fill.size() + n->elem_size, 0);
            append_message(msgs, 0x0005, fill);
            n->index_field = n->header_address + 16 + msgs.size() + 8 + 3;  // B-tree address inside the layout message
            std::vector<uint8_t> lay = {3, 2, (uint8_t)(n->row_shape.size() + 2)};
            put<uint64_t>(lay, kUndefined);  // no chunk yet
            put<uint32_t>(lay, (uint32_t)n->chunk_rows);
            for (uint64_t d : n->row_shape) put<uint32_t>(lay, (uint32_t)d);
            put<uint32_t>(lay, n->elem_size);
            append_message(msgs, 0x0008, lay);
        } else {
            append_message(msgs, 0x0001, n->dataspace);
            append_message(msgs, 0x0003, n->datatype, 1);
            append_message(msgs, 0x0005, {2, 2, 2, 0, 0, 0, 0, 0});  // fill value v2: allocate late, write if set, undefined
            std::vector<uint8_t> lay = {3, 1};
            put<uint64_t>(lay, n->data_address);
            put<uint64_t>(lay, (uint64_t)n->owned.size());
            append_message(msgs, 0x0008, lay);
        }
        for (const auto& a : n->attribute_messages) append_message(msgs, 0x000C, a.second);
        std::vector<uint8_t> hdr = {1, 0};
        put<uint16_t>(hdr, (uint16_t)((n->is_group ? 1 : 4) + n->attribute_messages.size()));
        put<uint32_t>(hdr, 1);  // reference count
        put<uint32_t>(hdr, (uint32_t)msgs.size());
        put<uint32_t>(hdr, 0);  // alignment of the messages to 8 bytes
        store(n->header_address, hdr);
        store(n->header_address + 16, msgs);

        if (!n->is_group) {
            if (!n->extendible) store(n->data_address, n->owned);
            continue;
        }
        // local heap: names sorted (std::map order == strcmp order for plain ASCII)
        std::vector<uint8_t> heap_data(n->data_bytes, 0);
        std::vector<std::pair<uint64_t, Node*>> entries;
        uint64_t off = 8;
        uint64_t last_name_off = 0;
        for (const auto& kv : n->children) {
            std::memcpy(&heap_data[off], kv.first.c_str(), kv.first.size() + 1);
            entries.emplace_back(off, kv.second);
            last_name_off = off;
            off += pad8(kv.first.size() + 1);
        }
        const uint64_t free_off = off;  // one free block: next-free offset (1 = none), size
        uint64_t one = 1, free_size = n->data_bytes - free_off;
        std::memcpy(&heap_data[free_off], &one, 8);
        std::memcpy(&heap_data[free_off + 8], &free_size, 8);
        std::vector<uint8_t> heap;
        put_bytes(heap, "HEAP", 4);
        heap.insert(heap.end(), {0, 0, 0, 0});
        put<uint64_t>(heap, n->data_bytes);
        put<uint64_t>(heap, free_off);
        put<uint64_t>(heap, n->heap_data);
        store(n->heap, heap);
        store(n->heap_data, heap_data);

        // B-tree node type 0, level 0, one child (the symbol node); keys: "" and the largest name
        std::vector<uint8_t> bt;
        put_bytes(bt, "TREE", 4);
        bt.push_back(0);
        bt.push_back(0);
        put<uint16_t>(bt, (uint16_t)(entries.empty() ? 0 : 1));
        put<uint64_t>(bt, kUndefined);
        put<uint64_t>(bt, kUndefined);
        put<uint64_t>(bt, 0);
        put<uint64_t>(bt, n->snod);
        put<uint64_t>(bt, last_name_off);
        store(n->btree, bt);

        std::vector<uint8_t> sn;
        put_bytes(sn, "SNOD", 4);
        sn.push_back(1);
        sn.push_back(0);
        put<uint16_t>(sn, (uint16_t)entries.size());
        for (const auto& e : entries) {
            put<uint64_t>(sn, e.first);
            put<uint64_t>(sn, e.second->header_address);
            if (e.second->is_group) {
                put<uint32_t>(sn, 1);
                put<uint32_t>(sn, 0);
                put<uint64_t>(sn, e.second->btree);
                put<uint64_t>(sn, e.second->heap);
            } else {
                put<uint32_t>(sn, 0);
                put<uint32_t>(sn, 0);
                put<uint64_t>(sn, 0);
                put<uint64_t>(sn, 0);
            }
        }
        store(n->snod, sn);
    }

    FILE* out = std::fopen(filename.c_str(), "wb");
    if (!out) throw Error("cannot open " + filename + " for writing");
    const bool ok = std::fwrite(f.data(), 1, f.size(), out) == f.size();
    std::fclose(out);
    if (!ok) throw Error("short write on " + filename);
    // rows appended before the file existed
    for (Node* n : extendibles) {
        if (n->pending.empty()) continue;
        std::vector<uint8_t> rows;
        rows.swap(n->pending);
        append_to(n, rows.data(), rows.size() / n->row_bytes);
    }
    flush();
}

}  // namespace h5lite
