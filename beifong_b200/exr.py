"""Minimal multichannel OpenEXR scanline files (uncompressed, HALF or FLOAT channels): what `bitmap()->write(filename, OpenEXR)` of
HDRADC::develop() produces (src/adcs/hdradc.cpp:277-295; src/libcore/bitmap.cpp write_openexr) for a MultiChannel bitmap -- one EXR
channel per Struct field, channel names = field names, stored in alphabetical order as the format requires.  The reference links
the OpenEXR library; this writer emits the same container directly (file layout: OpenEXR "Technical Introduction" / file layout
document: magic 20000630, version 2, attribute list, line-offset table, one chunk per scan line for NO_COMPRESSION).
`read_exr` reads back what `write_exr` wrote (and any other uncompressed scanline file) for the tests."""
import struct

import numpy as np

MAGIC = 20000630
PIXEL_TYPES = {np.dtype(np.uint32): 0, np.dtype(np.float16): 1, np.dtype(np.float32): 2}
DTYPES = {0: np.dtype("<u4"), 1: np.dtype("<f2"), 2: np.dtype("<f4")}


def _attr(name, type_name, payload):
    return name.encode() + b"\0" + type_name.encode() + b"\0" + struct.pack("<i", len(payload)) + payload


def write_exr(path, image, channel_names):
    """image: [height][width][channels] float16 / float32 / uint32 array; channel_names: one name per channel."""
    image = np.asarray(image)
    if image.ndim != 3 or image.shape[2] != len(channel_names):
        raise ValueError("write_exr: image must be [h][w][len(channel_names)]")
    if image.dtype not in PIXEL_TYPES:
        raise ValueError(f"write_exr: unsupported component type {image.dtype}")
    if len(set(channel_names)) != len(channel_names):
        raise ValueError("write_exr: duplicate channel names")
    h, w, _ = image.shape
    ptype = PIXEL_TYPES[image.dtype]
    order = sorted(range(len(channel_names)), key=lambda i: channel_names[i].encode())
    chlist = b"".join(channel_names[i].encode() + b"\0" + struct.pack("<iB3xii", ptype, 0, 1, 1) for i in order) + b"\0"
    box = struct.pack("<iiii", 0, 0, w - 1, h - 1)
    header = struct.pack("<ii", MAGIC, 2)
    header += _attr("channels", "chlist", chlist)
    header += _attr("compression", "compression", b"\0")
    header += _attr("dataWindow", "box2i", box)
    header += _attr("displayWindow", "box2i", box)
    header += _attr("lineOrder", "lineOrder", b"\0")
    header += _attr("pixelAspectRatio", "float", struct.pack("<f", 1.0))
    header += _attr("screenWindowCenter", "v2f", struct.pack("<ff", 0.0, 0.0))
    header += _attr("screenWindowWidth", "float", struct.pack("<f", 1.0))
    header += b"\0"
    row_bytes = w * image.dtype.itemsize * len(order)
    first = len(header) + 8 * h
    offsets = struct.pack(f"<{h}Q", *[first + y * (8 + row_bytes) for y in range(h)])
    planar = np.ascontiguousarray(image[:, :, order].transpose(0, 2, 1)).astype(image.dtype.newbyteorder("<"))   # [y][channel][x]
    with open(path, "wb") as f:
        f.write(header)
        f.write(offsets)
        for y in range(h):
            f.write(struct.pack("<ii", y, row_bytes))
            f.write(planar[y].tobytes())
    return path


def read_exr(path):
    """-> (image [h][w][c], channel names in file order).  Uncompressed scanline files with one pixel type only."""
    data = open(path, "rb").read()
    magic, version = struct.unpack_from("<ii", data, 0)
    if magic != MAGIC or (version & 0xff) != 2 or (version & 0x200):
        raise ValueError("read_exr: not a single-part scanline OpenEXR 2 file")
    pos, attrs = 8, {}
    while data[pos] != 0:
        e = data.index(b"\0", pos); name = data[pos:e].decode(); pos = e + 1
        e = data.index(b"\0", pos); tname = data[pos:e].decode(); pos = e + 1
        (size,) = struct.unpack_from("<i", data, pos); pos += 4
        attrs[name] = (tname, data[pos:pos + size]); pos += size
    pos += 1
    if attrs["compression"][1] != b"\0":
        raise ValueError("read_exr: compressed files are not supported")
    names, ptypes, cl, p = [], [], attrs["channels"][1], 0
    while cl[p] != 0:
        e = cl.index(b"\0", p); names.append(cl[p:e].decode()); p = e + 1
        pt, _, xs, ys = struct.unpack_from("<iB3xii", cl, p); p += 16
        ptypes.append(pt)
    if len(set(ptypes)) != 1:
        raise ValueError("read_exr: mixed pixel types")
    x0, y0, x1, y1 = struct.unpack("<iiii", attrs["dataWindow"][1])
    w, h = x1 - x0 + 1, y1 - y0 + 1
    dt = DTYPES[ptypes[0]]
    offsets = struct.unpack_from(f"<{h}Q", data, pos)
    out = np.empty((h, len(names), w), dtype=dt)
    for off in offsets:
        y, size = struct.unpack_from("<ii", data, off)
        out[y - y0] = np.frombuffer(data, dtype=dt, count=len(names) * w, offset=off + 8).reshape(len(names), w)
    return np.ascontiguousarray(out.transpose(0, 2, 1)), names
