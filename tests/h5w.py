"""Test tooling for HDF5 files: (1) ctypes over libh5lite.so (h5lite::Writer) to compose files in the reference's
schemas, (2) a tiny independent classic-format writer in pure Python for the reader features the C++ writer does not
produce (chunked + deflate + shuffle datasets, variable-length string attributes in a global heap), (3) h5ls() to read
files back through the C++ reader."""
import ctypes as C
import json
import os
import struct
import subprocess
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "mpi-cuda-sartsolver_b200", "host")
_lib = None


def _capi():
    global _lib
    if _lib is None:
        so = os.path.join(HOST, "libh5lite.so")
        if not os.path.exists(so):
            subprocess.check_call(["make", "-C", HOST, "libh5lite.so", "sart_b200_h5ls"])
        L = C.CDLL(so)
        L.h5w_open.restype = C.c_void_p
        L.h5w_error.restype = C.c_char_p
        for name in ("h5w_f64", "h5w_f32", "h5w_i32", "h5w_u64"):
            getattr(L, name).argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int, C.c_void_p]
        L.h5w_group.argtypes = [C.c_void_p, C.c_char_p]
        L.h5w_attr_f64.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_double]
        L.h5w_attr_u64.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_uint64]
        L.h5w_attr_i32.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int32]
        L.h5w_attr_str.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p]
        L.h5w_close.argtypes = [C.c_void_p]
        L.h5w_extendible.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_void_p, C.c_int, C.c_uint64]
        L.h5w_append.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_uint64]
        L.h5w_layout.argtypes = [C.c_void_p]
        L.h5w_flush.argtypes = [C.c_void_p]
        L.h5w_open.argtypes = [C.c_char_p]
        _lib = L
    return _lib


class H5Writer:
    def __init__(self, path):
        self.L = _capi()
        self.h = self.L.h5w_open(os.fsencode(path))
        assert self.h, self.L.h5w_error()
        self._keep = []

    def _ok(self, rc):
        assert rc == 0, self.L.h5w_error().decode()

    def group(self, path):
        self._ok(self.L.h5w_group(self.h, path.encode()))

    def dataset(self, path, array):
        a = np.ascontiguousarray(array)
        fn = {np.dtype("float64"): self.L.h5w_f64, np.dtype("float32"): self.L.h5w_f32, np.dtype("int32"): self.L.h5w_i32,
              np.dtype("uint64"): self.L.h5w_u64}[a.dtype]
        dims = (C.c_uint64 * max(a.ndim, 1))(*a.shape)
        self._ok(fn(self.h, path.encode(), dims, a.ndim, a.ctypes.data_as(C.c_void_p)))

    def attr(self, obj, name, value):
        if isinstance(value, str):
            self._ok(self.L.h5w_attr_str(self.h, obj.encode(), name.encode(), value.encode()))
        elif isinstance(value, float):
            self._ok(self.L.h5w_attr_f64(self.h, obj.encode(), name.encode(), value))
        elif isinstance(value, np.int32):
            self._ok(self.L.h5w_attr_i32(self.h, obj.encode(), name.encode(), int(value)))
        else:
            self._ok(self.L.h5w_attr_u64(self.h, obj.encode(), name.encode(), int(value)))

    def extendible(self, path, dtype, row_shape, chunk_rows):
        """Dataset with an unlimited first dimension (float64 or int32 rows of ``row_shape``), chunks of ``chunk_rows`` rows."""
        shape = (C.c_uint64 * max(len(row_shape), 1))(*row_shape)
        self._ok(self.L.h5w_extendible(self.h, path.encode(), int(np.dtype(dtype) == np.float64), shape, len(row_shape), chunk_rows))

    def append(self, path, rows):
        a = np.ascontiguousarray(rows)
        assert a.dtype in (np.float64, np.int32)
        self._ok(self.L.h5w_append(self.h, path.encode(), a.ctypes.data_as(C.c_void_p), a.shape[0]))

    def layout(self):
        """Writes the file; extendible datasets stay open for append() / flush()."""
        self._ok(self.L.h5w_layout(self.h))

    def flush(self):
        self._ok(self.L.h5w_flush(self.h))

    def close(self):
        self._ok(self.L.h5w_close(self.h))
        self.h = None


def h5ls(path, values=True):
    exe = os.path.join(HOST, "sart_b200_h5ls")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", HOST, "sart_b200_h5ls"])
    out = subprocess.run([exe, str(path)] + (["--values"] if values else []), capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    return {o["path"]: o for o in (json.loads(l) for l in out.stdout.splitlines())}


# ---- the reference's schemas ---------------------------------------------------------------------

def write_rtm_h5(path, block, camera, rtm_name, mask, voxel_cells, grid, wavelength=656.3, sparse=False):
    """One camera x voxel-segment file: rtm attrs, rtm/<name>/value (dense) or triplets (sparse), voxel_map, frame_mask."""
    w = H5Writer(path)
    w.group("rtm")
    w.attr("rtm", "camera_name", camera)
    w.attr("rtm", "npixel", block.shape[0])
    w.attr("rtm", "nvoxel", block.shape[1])
    w.group("rtm/" + rtm_name)
    w.attr("rtm/" + rtm_name, "wavelength", float(wavelength))
    w.attr("rtm/" + rtm_name, "is_sparse", np.int32(1 if sparse else 0))
    if sparse:
        pi, vi = np.nonzero(block)
        w.dataset(f"rtm/{rtm_name}/pixel_index", pi.astype(np.uint64))
        w.dataset(f"rtm/{rtm_name}/voxel_index", vi.astype(np.uint64))
        w.dataset(f"rtm/{rtm_name}/value", block[pi, vi].astype(np.float32))
    else:
        w.dataset(f"rtm/{rtm_name}/value", block.astype(np.float32))
    nx, ny, nz = grid
    w.group("rtm/voxel_map")
    for n, v in zip(("nx", "ny", "nz"), grid):
        w.attr("rtm/voxel_map", n, v)
    w.attr("rtm/voxel_map", "xmin", 1.5)
    w.attr("rtm/voxel_map", "xmax", 4.5)
    w.attr("rtm/voxel_map", "coordinate_system", "Cylindrical")
    cells = np.asarray(voxel_cells, dtype=np.uint64)
    w.dataset("rtm/voxel_map/i", cells // np.uint64(ny * nz))
    w.dataset("rtm/voxel_map/j", (cells % np.uint64(ny * nz)) // np.uint64(nz))
    w.dataset("rtm/voxel_map/k", cells % np.uint64(nz))
    w.dataset("rtm/voxel_map/value", np.arange(len(cells), dtype=np.int32))
    w.dataset("rtm/frame_mask", np.asarray(mask, dtype=np.int32))
    w.close()


def write_image_h5(path, camera, times, frames3d, wavelength=656.3):
    w = H5Writer(path)
    w.group("image")
    w.attr("image", "camera_name", camera)
    w.attr("image", "wavelength", float(wavelength))
    w.dataset("image/time", np.asarray(times, dtype=np.float64))
    w.dataset("image/frame", np.asarray(frames3d, dtype=np.float64))
    w.close()


def write_laplacian_h5(path, flat_index, value, nvoxel):
    flat_index = np.asarray(flat_index, dtype=np.uint64)
    w = H5Writer(path)
    w.group("laplacian")
    w.attr("laplacian", "nvoxel", nvoxel)
    w.dataset("laplacian/i", flat_index // np.uint64(nvoxel))
    w.dataset("laplacian/j", flat_index % np.uint64(nvoxel))
    w.dataset("laplacian/value", np.asarray(value, dtype=np.float32))
    w.close()


# ---- independent mini writer (pure Python): chunked/deflate/shuffle dataset + vlen string attribute ---------------------

def _pad8(b):
    return b + b"\0" * (-len(b) % 8)


def _msg(mtype, body, flags=0):
    body = _pad8(body)
    return struct.pack("<HHB3x", mtype, len(body), flags) + body


def write_chunked_h5(path, array, chunk, deflate=True, shuffle=True, vlen_attr=("camera_name", "cam_vlen")):
    """Classic-format file with root group -> dataset "data" (2-D float32, chunked, v1 B-tree) and a variable-length string
    attribute on the root group. Written from the format description, independently of h5lite's writer."""
    a = np.ascontiguousarray(array, dtype=np.float32)
    rows, cols = a.shape
    cr, cc = chunk
    UNDEF = 0xFFFFFFFFFFFFFFFF
    blob = bytearray(b"\0" * 96)

    def alloc(data):
        while len(blob) % 8:
            blob.append(0)
        addr = len(blob)
        blob.extend(data)
        return addr

    # chunks
    keys = []
    for r0 in range(0, rows, cr):
        for c0 in range(0, cols, cc):
            tile = np.zeros((cr, cc), dtype=np.float32)
            sub = a[r0:r0 + cr, c0:c0 + cc]
            tile[:sub.shape[0], :sub.shape[1]] = sub
            raw = tile.tobytes()
            if shuffle:
                raw = np.frombuffer(raw, np.uint8).reshape(-1, 4).T.tobytes()
            if deflate:
                raw = zlib.compress(raw, 6)
            keys.append((len(raw), r0, c0, alloc(raw)))
    # chunk B-tree: one leaf node (type 1, level 0)
    node = b"TREE" + struct.pack("<BBH", 1, 0, len(keys)) + struct.pack("<QQ", UNDEF, UNDEF)
    for size, r0, c0, addr in keys:
        node += struct.pack("<II3Q", size, 0, r0, c0, 0) + struct.pack("<Q", addr)
    node += struct.pack("<II3Q", 0, 0, rows, cols, 0)  # final key
    btree = alloc(node)

    # global heap with the vlen string
    text = vlen_attr[1].encode()
    gcol_body = struct.pack("<HHIQ", 1, 1, 0, len(text)) + _pad8(text)
    gcol_size = 16 + len(gcol_body) + 16
    gcol = b"GCOL" + struct.pack("<B3xQ", 1, gcol_size) + gcol_body + struct.pack("<HHIQ", 0, 0, 0, 16)
    gcol_addr = alloc(gcol)

    # dataset object header
    dtype_f32 = bytes([0x11, 0x20, 31, 0]) + struct.pack("<IHH4BI", 4, 0, 32, 23, 8, 0, 23, 127)
    space = struct.pack("<BBB5x", 1, 2, 0) + struct.pack("<QQ", rows, cols)
    layout = struct.pack("<BBB", 3, 2, 3) + struct.pack("<Q", btree) + struct.pack("<III", cr, cc, 4)
    filters = b""
    nf = 0
    if shuffle:
        filters += struct.pack("<HHHH", 2, 0, 0, 1) + struct.pack("<I", 4) + b"\0" * 4
        nf += 1
    if deflate:
        filters += struct.pack("<HHHH", 1, 0, 0, 1) + struct.pack("<I", 6) + b"\0" * 4
        nf += 1
    msgs = _msg(1, space) + _msg(3, dtype_f32, 1) + _msg(8, layout)
    if nf:
        msgs += _msg(0xB, struct.pack("<BB6x", 1, nf) + filters)
    dset_hdr = alloc(struct.pack("<BBHII4x", 1, 0, 3 + (1 if nf else 0), 1, len(msgs)) + msgs)

    # root group: heap, symbol node, B-tree, header with the vlen attribute
    heap_data = b"\0" * 8 + _pad8(b"data\0") + struct.pack("<QQ", 1, 16)
    heap_data_addr = alloc(heap_data)
    heap = alloc(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), 16, heap_data_addr))
    snod = alloc(b"SNOD" + struct.pack("<BBH", 1, 0, 1) + struct.pack("<QQII16x", 8, dset_hdr, 0, 0) + b"\0" * (40 * 7))
    gtree = alloc(b"TREE" + struct.pack("<BBH", 0, 0, 1) + struct.pack("<QQ", UNDEF, UNDEF) + struct.pack("<QQQ", 0, snod, 8) + b"\0" * 512)
    vlen_type = bytes([0x19, 0x01, 0, 0]) + struct.pack("<I", 16) + bytes([0x10, 0, 0, 0]) + struct.pack("<IHH", 1, 0, 8)
    name = vlen_attr[0].encode() + b"\0"
    attr = struct.pack("<BBHHH", 1, 0, len(name), len(vlen_type), 8) + _pad8(name) + _pad8(vlen_type) + struct.pack("<BBB5x", 1, 0, 0)
    attr += struct.pack("<IQI", len(text), gcol_addr, 1)
    rmsgs = _msg(0x11, struct.pack("<QQ", gtree, heap)) + _msg(0xC, attr)
    root_hdr = alloc(struct.pack("<BBHII4x", 1, 0, 2, 1, len(rmsgs)) + rmsgs)

    sb = bytes([0x89]) + b"HDF\r\n\x1a\n" + bytes([0, 0, 0, 0, 0, 8, 8, 0]) + struct.pack("<HHI", 4, 16, 0)
    sb += struct.pack("<QQQQ", 0, UNDEF, len(blob), UNDEF) + struct.pack("<QQII", 0, root_hdr, 1, 0) + struct.pack("<QQ", gtree, heap)
    blob[0:len(sb)] = sb
    with open(path, "wb") as f:
        f.write(bytes(blob))


# ---- a "libver latest" style file, written from the format description -----------------------------------------------------

def _msg2(mtype, body, flags=0):
    """Version-2 object header message: type u8, size u16, flags u8, body (not padded)."""
    return struct.pack("<BHB", mtype, len(body), flags) + body


def _ohdr(messages, extra_block=None):
    """"OHDR" header with one chunk; ``extra_block`` = (address, length) of a continuation block to point at."""
    body = b"".join(messages)
    if extra_block is not None:
        body += _msg2(0x10, struct.pack("<QQ", *extra_block))
    # flags 0x01: chunk-0 size in two bytes; checksum left zero (readers here do not verify it)
    return b"OHDR" + bytes([2, 0x01]) + struct.pack("<H", len(body)) + body + b"\0\0\0\0"


def write_latest_h5(path, a, attr_value=2.5):
    """Superblock 2 -> root group (link messages) -> group "g" -> the float64 matrix ``a`` stored five ways under layout
    version 4: contiguous, compact, a single chunk, implicit chunks, fixed-array-indexed deflated chunks. One dataset's
    header continues in an "OCHK" block. For the reader features that no writer of this repository produces."""
    a = np.ascontiguousarray(a, dtype=np.float64)
    rows, cols = a.shape
    UNDEF = 0xFFFFFFFFFFFFFFFF
    blob = bytearray(b"\0" * 48)

    def alloc(data):
        while len(blob) % 8:
            blob.append(0)
        addr = len(blob)
        blob.extend(data)
        return addr

    dtype_f64 = bytes([0x11, 0x20, 63, 0]) + struct.pack("<IHH4BI", 8, 0, 64, 52, 11, 0, 52, 1023)
    space = struct.pack("<BBBB", 2, 2, 0, 1) + struct.pack("<QQ", rows, cols)  # dataspace version 2, simple
    common = [_msg2(1, space), _msg2(3, dtype_f64, 1)]
    links = {}

    # contiguous
    addr = alloc(a.tobytes())
    links["contiguous"] = alloc(_ohdr(common + [_msg2(8, struct.pack("<BBQQ", 4, 1, addr, a.nbytes))]))
    # compact (data inside the layout message)
    small = a[:2, :3].copy()
    sp_small = struct.pack("<BBBB", 2, 2, 0, 1) + struct.pack("<QQ", 2, 3)
    links["compact"] = alloc(_ohdr([_msg2(1, sp_small), _msg2(3, dtype_f64, 1),
                                    _msg2(8, struct.pack("<BBH", 4, 0, small.nbytes) + small.tobytes())]))
    # a single chunk as large as the dataset; its header continues in a second block that holds an attribute
    addr = alloc(a.tobytes())
    lay = struct.pack("<BBBBB", 4, 2, 0, 3, 4) + struct.pack("<III", rows, cols, 8) + bytes([1]) + struct.pack("<Q", addr)
    name = b"scale\0"
    attr = struct.pack("<BBHHHB", 3, 0, len(name), len(dtype_f64), 4, 0) + name + dtype_f64 + struct.pack("<BBBB", 2, 0, 0, 0)
    attr += struct.pack("<d", attr_value)
    ochk = b"OCHK" + _msg2(0xC, attr) + b"\0\0\0\0"
    ochk_addr = alloc(ochk)
    links["single"] = alloc(_ohdr(common + [_msg2(8, lay)], extra_block=(ochk_addr, len(ochk))))
    # implicit index: all chunks allocated in chunk-grid order from one address, no filters
    cr, cc = max(1, rows // 3), max(1, cols // 2)
    tiles = []
    for r0 in range(0, rows, cr):
        for c0 in range(0, cols, cc):
            tile = np.zeros((cr, cc))
            sub = a[r0:r0 + cr, c0:c0 + cc]
            tile[:sub.shape[0], :sub.shape[1]] = sub
            tiles.append(tile)
    addr = alloc(b"".join(t.tobytes() for t in tiles))
    lay = struct.pack("<BBBBB", 4, 2, 0, 3, 4) + struct.pack("<III", cr, cc, 8) + bytes([2]) + struct.pack("<Q", addr)
    links["implicit"] = alloc(_ohdr(common + [_msg2(8, lay)]))
    # fixed-array index over deflated chunks (client 1: address, stored size, filter mask per chunk)
    elems = b""
    for t in tiles:
        raw = zlib.compress(t.tobytes(), 6)
        elems += struct.pack("<QII", alloc(raw), len(raw), 0)
    fahd_addr = alloc(b"\0" * 28)
    fadb_addr = alloc(b"FADB" + bytes([0, 1]) + struct.pack("<Q", fahd_addr) + elems + b"\0\0\0\0")
    blob[fahd_addr:fahd_addr + 28] = b"FAHD" + bytes([0, 1, 16, 10]) + struct.pack("<QQ", len(tiles), fadb_addr) + b"\0\0\0\0"
    lay = struct.pack("<BBBBB", 4, 2, 0, 3, 4) + struct.pack("<III", cr, cc, 8) + bytes([3, 10]) + struct.pack("<Q", fahd_addr)
    pipeline = struct.pack("<BB", 2, 1) + struct.pack("<HHH", 1, 0, 1) + struct.pack("<I", 6)
    links["fixed_array"] = alloc(_ohdr(common + [_msg2(8, lay), _msg2(0xB, pipeline)]))

    def link(name, target):  # hard link, one-byte name length, no optional fields
        n = name.encode()
        return _msg2(6, struct.pack("<BBB", 1, 0, len(n)) + n + struct.pack("<Q", target))

    link_info = _msg2(2, struct.pack("<BBQQ", 0, 0, UNDEF, UNDEF))  # compact storage: no fractal heap
    group = alloc(_ohdr([link_info] + [link(k, v) for k, v in sorted(links.items())]))
    root = alloc(_ohdr([link_info, link("g", group)]))
    sb = bytes([0x89]) + b"HDF\r\n\x1a\n" + bytes([2, 8, 8, 0]) + struct.pack("<QQQQ", 0, UNDEF, len(blob), root) + b"\0\0\0\0"
    blob[0:len(sb)] = sb
    with open(path, "wb") as f:
        f.write(bytes(blob))
    return small
