"""Headless image output for the developed film: what `HDRFilm::write` / `mi.util.write_bitmap` do for this path
(src/films/hdrfilm.cpp:546-560 writes the developed bitmap as OpenEXR; src/python/python/util.py write_bitmap).

Only what the render path needs: float32 RGB / RGBA / luminance images written as uncompressed scanline OpenEXR (readable
by any EXR viewer) or PFM, and readers for the files written here (used by the tests for round trips). No external
library: the OpenEXR 2.0 single-part scanline layout is a magic number, a version word, a list of named attributes, a
table of scanline offsets and, per scanline, `y, byte count, channel planes` with the channels in alphabetical order.
"""
import struct

import numpy as np

_EXR_MAGIC = 20000630
_CHANNELS = {1: ("Y",), 3: ("R", "G", "B"), 4: ("R", "G", "B", "A")}


def _to_host(img) -> np.ndarray:
    if hasattr(img, "detach"):
        img = img.detach().cpu().numpy()
    a = np.asarray(img)
    if a.ndim == 2:
        a = a[..., None]
    if a.ndim != 3 or a.shape[2] not in _CHANNELS:
        raise ValueError(f"expected an [H,W], [H,W,1], [H,W,3] or [H,W,4] image, got shape {a.shape}")
    return a


def _attr(name: str, typ: str, payload: bytes) -> bytes:
    return name.encode() + b"\0" + typ.encode() + b"\0" + struct.pack("<i", len(payload)) + payload


def write_exr(path, img, half=False):
    """Write an [H,W,C] float image (C = 1, 3, 4) as an uncompressed scanline OpenEXR file (float32, or float16 if half)."""
    a = _to_host(img).astype(np.float16 if half else np.float32)
    h, w, c = a.shape
    names = _CHANNELS[c]
    order = sorted(range(c), key=lambda i: names[i])            # planes are stored in alphabetical channel order
    pixel_type = 1 if half else 2                               # 1 = HALF, 2 = FLOAT
    chlist = b"".join(names[i].encode() + b"\0" + struct.pack("<iBxxxii", pixel_type, 0, 1, 1) for i in order) + b"\0"
    box = struct.pack("<iiii", 0, 0, w - 1, h - 1)
    header = (_attr("channels", "chlist", chlist) + _attr("compression", "compression", b"\0") +
              _attr("dataWindow", "box2i", box) + _attr("displayWindow", "box2i", box) +
              _attr("lineOrder", "lineOrder", b"\0") + _attr("pixelAspectRatio", "float", struct.pack("<f", 1.0)) +
              _attr("screenWindowCenter", "v2f", struct.pack("<ff", 0.0, 0.0)) +
              _attr("screenWindowWidth", "float", struct.pack("<f", 1.0)) + b"\0")
    row_bytes = w * c * a.itemsize
    start = 8 + len(header) + 8 * h
    offsets = struct.pack(f"<{h}Q", *[start + y * (8 + row_bytes) for y in range(h)])
    planes = np.ascontiguousarray(a[:, :, order].transpose(0, 2, 1))   # [H, C(alphabetical), W]
    with open(path, "wb") as f:
        f.write(struct.pack("<ii", _EXR_MAGIC, 2))
        f.write(header)
        f.write(offsets)
        for y in range(h):
            f.write(struct.pack("<ii", y, row_bytes))
            f.write(planes[y].tobytes())


def read_exr(path) -> np.ndarray:
    """Read an uncompressed scanline OpenEXR file as written by write_exr -> [H,W,C] float32 (channels R,G,B[,A] or Y)."""
    with open(path, "rb") as f:
        data = f.read()
    magic, version = struct.unpack_from("<ii", data, 0)
    if magic != _EXR_MAGIC or (version & 0xff) != 2 or (version & ~0xff):
        raise ValueError("not a single-part scanline OpenEXR 2.0 file")
    pos, attrs = 8, {}
    while data[pos] != 0:
        e = data.index(b"\0", pos); name = data[pos:e].decode(); pos = e + 1
        e = data.index(b"\0", pos); pos = e + 1
        (size,) = struct.unpack_from("<i", data, pos); pos += 4
        attrs[name] = data[pos:pos + size]; pos += size
    pos += 1
    if attrs["compression"] != b"\0":
        raise ValueError("only uncompressed files are supported")
    x0, y0, x1, y1 = struct.unpack("<iiii", attrs["dataWindow"])
    w, h = x1 - x0 + 1, y1 - y0 + 1
    names, types, p, cl = [], [], 0, attrs["channels"]
    while cl[p] != 0:
        e = cl.index(b"\0", p); names.append(cl[p:e].decode()); p = e + 1
        types.append(struct.unpack_from("<i", cl, p)[0]); p += 16
    if len(set(types)) != 1 or types[0] not in (1, 2):
        raise ValueError("mixed or integer channel types are not supported")
    dt = np.float16 if types[0] == 1 else np.float32
    offsets = struct.unpack_from(f"<{h}Q", data, pos)
    out = np.zeros((h, w, len(names)), np.float32)
    want = _CHANNELS[len(names)]
    for off in offsets:
        y, nbytes = struct.unpack_from("<ii", data, off)
        row = np.frombuffer(data, dt, w * len(names), off + 8).reshape(len(names), w)
        for k, nm in enumerate(names):
            out[y - y0, :, want.index(nm)] = row[k]
    return out


def write_pfm(path, img):
    """Portable float map (bottom-to-top scanlines, little endian): [H,W,1] or [H,W,3] float32."""
    a = _to_host(img).astype(np.float32)
    if a.shape[2] == 4:
        a = a[..., :3]
    h, w, c = a.shape
    with open(path, "wb") as f:
        f.write(b"PF\n" if c == 3 else b"Pf\n")
        f.write(f"{w} {h}\n-1.0\n".encode())
        f.write(np.ascontiguousarray(a[::-1]).tobytes())


def read_pfm(path) -> np.ndarray:
    with open(path, "rb") as f:
        kind = f.readline().strip()
        w, h = (int(v) for v in f.readline().split())
        scale = float(f.readline())
        c = 3 if kind == b"PF" else 1
        a = np.frombuffer(f.read(), "<f4" if scale < 0 else ">f4", w * h * c).reshape(h, w, c)
    return np.ascontiguousarray(a[::-1]).astype(np.float32)


def write_bitmap(path, img, **kw):
    """mi.util.write_bitmap for the formats of this build: .exr (default) and .pfm, chosen by the file extension."""
    p = str(path).lower()
    if p.endswith(".pfm"):
        return write_pfm(path, img)
    if p.endswith(".exr"):
        return write_exr(path, img, **kw)
    raise ValueError("write_bitmap: only .exr and .pfm are supported in this build (no PNG/JPEG encoders)")
