"""The on-disk contract of the path's output: the raw `-f` frame
must be readable however a TIFF library lays it out.  libtiff -- which the
reference engine writes through and its post-processor reads with
(raytracer/image.c:64-139, postprocess/image.c:30-73) -- is not in this image,
so the layouts it produces are synthesised here: the IFD after the pixel data,
strips of several rows, SHORT strip byte counts, additional ASCII / RATIONAL /
SampleFormat tags, either byte order.  The reader is src/raytracer/tiff.c
(rt_tiff_read through host.load_raw), the one the `postprocess` executable uses."""
import os
import struct
import subprocess

import numpy as np
import pytest

from _util import ROOT
from c_raytracer_b200 import host


def synth_tiff(path, raster, z, order="<", rows_per_strip=1, ifd_last=True, short_counts=False, extras=True, bits=32):
    """A classic TIFF the way a general-purpose writer lays it out."""
    h, w = raster.shape[:2]
    e = order
    if bits == 32:
        pix = raster.astype(e + "f4").tobytes()
    else:
        pix = raster.astype(np.uint8).tobytes()
    row_bytes = w * 3 * bits // 8
    strips = [pix[r * row_bytes:(r + rows_per_strip) * row_bytes] for r in range(0, h, rows_per_strip)]
    blob = bytearray(b"II*\x00" if e == "<" else b"MM\x00*")
    blob += b"\x00" * 4  # IFD offset, patched below
    extra_data = {}

    def put(data):
        if len(blob) % 2:
            blob.append(0)
        off = len(blob)
        blob.extend(data)
        return off

    strip_off = [put(s) for s in strips] if ifd_last else None
    entries = []  # (tag, type, count, value bytes or offset)

    def entry(tag, typ, values):
        size = {1: 1, 2: 1, 3: 2, 4: 4, 5: 8, 11: 4}[typ]
        fmt = {1: "B", 2: "s", 3: "H", 4: "I", 11: "f"}.get(typ)
        if typ == 2:
            data = values
            count = len(values)
        elif typ == 5:
            data = b"".join(struct.pack(e + "II", *v) for v in values)
            count = len(values)
        else:
            data = struct.pack(e + str(len(values)) + fmt, *values)
            count = len(values)
        entries.append((tag, typ, count, data if size * count <= 4 else None, data))

    entry(256, 4, [w])
    entry(257, 4, [h])
    entry(258, 3, [bits] * 3)
    entry(259, 3, [1])
    entry(262, 3, [2])
    if extras:
        entry(305, 2, b"synthetic writer 1.0\x00")
        entry(282, 5, [(72, 1)])
        entry(283, 5, [(72, 1)])
        entry(296, 3, [2])
        if bits == 32:
            entry(339, 3, [3] * 3)  # SampleFormat = IEEE float: what a careful writer adds
    entry(274, 3, [1])
    entry(277, 3, [3])
    entry(278, 4, [rows_per_strip])
    entry(284, 3, [1])
    if z is not None:
        entry(65000, 11, list(z.astype("f4").reshape(-1)))
    counts = [len(s) for s in strips]
    entries_sorted = None
    # out-of-line data first, then strips (if not already written), then the IFD
    pending = []
    for tag, typ, count, inline, data in entries:
        pending.append([tag, typ, count, inline, data])
    if not ifd_last:
        ifd_off = put(b"")  # IFD right after the header
        n = len(pending) + 2
        blob.extend(b"\x00" * (2 + 12 * n + 4))
    for p in pending:
        if p[3] is None:
            p[3] = struct.pack(e + "I", put(p[4]))
        else:
            p[3] = p[3].ljust(4, b"\x00")
    if strip_off is None:
        strip_off = [put(s) for s in strips]
    off_data = struct.pack(e + str(len(strip_off)) + "I", *strip_off)
    cnt_fmt, cnt_type = ("H", 3) if short_counts and max(counts) < 65536 else ("I", 4)
    cnt_data = struct.pack(e + str(len(counts)) + cnt_fmt, *counts)
    size = {3: 2, 4: 4}
    pending.append([273, 4, len(strip_off), off_data.ljust(4, b"\x00") if 4 * len(strip_off) <= 4 else struct.pack(e + "I", put(off_data)), off_data])
    pending.append([279, cnt_type, len(counts),
                    cnt_data.ljust(4, b"\x00") if size[cnt_type] * len(counts) <= 4 else struct.pack(e + "I", put(cnt_data)), cnt_data])
    pending.sort(key=lambda p: p[0])
    ifd = struct.pack(e + "H", len(pending)) + b"".join(struct.pack(e + "HHI", p[0], p[1], p[2]) + p[3] for p in pending) + b"\x00" * 4
    if ifd_last:
        ifd_off = put(ifd)
    else:
        blob[ifd_off:ifd_off + len(ifd)] = ifd
    blob[4:8] = struct.pack(e + "I", ifd_off)
    with open(path, "wb") as f:
        f.write(blob)


@pytest.mark.parametrize("order", ["<", ">"])
@pytest.mark.parametrize("rows_per_strip,ifd_last,short_counts,extras", [
    (1, True, False, True), (1, False, True, False), (4, True, True, True), (1000, True, False, True), (3, False, False, True),
])
def test_raw_frame_in_library_layouts(tmp_path, order, rows_per_strip, ifd_last, short_counts, extras):
    rng = np.random.default_rng(5)
    h, w = 11, 9
    raster = rng.random((h, w, 3), dtype=np.float32)
    z = rng.random((h, w), dtype=np.float32)
    p = str(tmp_path / "lib.tif")
    synth_tiff(p, raster, z if order == "<" else None, order, min(rows_per_strip, h), ifd_last, short_counts, extras)
    if order == "<":
        r2, z2 = host.load_raw(p)
        np.testing.assert_array_equal(r2, raster)
        np.testing.assert_array_equal(z2, z)
    else:
        # tag 65000 is written by the engine in the machine's byte order only; a big-endian
        # frame without it is not a raw frame of this path
        with pytest.raises(host.SceneError, match="raw frame"):
            host.load_raw(p)


def test_pillow_reads_what_the_synthesiser_writes(tmp_path):
    """The synthesiser itself is checked against an independent TIFF reader."""
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(6)
    img = (rng.random((10, 7, 3)) * 255).astype(np.uint8)
    for order in "<>":
        p = str(tmp_path / f"p{ord(order)}.tif")
        synth_tiff(p, img, None, order, 3, True, True, True, bits=8)
        np.testing.assert_array_equal(np.array(Image.open(p)), img)


def test_truncated_and_foreign_files_are_rejected(tmp_path):
    rng = np.random.default_rng(7)
    raster = rng.random((6, 5, 3), dtype=np.float32)
    z = rng.random((6, 5), dtype=np.float32)
    p = str(tmp_path / "a.tif")
    host.save_tiff(p, raster, z, raw=True)
    b = open(p, "rb").read()
    for name, data in [("short.tif", b[:len(b) // 2]), ("empty.tif", b""), ("png.tif", b"\x89PNG\r\n\x1a\n" + b"\x00" * 64),
                       ("bigtiff.tif", b"II+\x00" + b"\x00" * 64)]:
        q = str(tmp_path / name)
        open(q, "wb").write(data)
        with pytest.raises(host.SceneError):
            host.load_raw(q)
    with pytest.raises(host.SceneError, match="open"):
        host.load_raw(str(tmp_path / "missing.tif"))


def _patch_ifd(blob, tag, value):
    """Overwrites the inline value of one IFD entry of a little-endian classic TIFF."""
    b = bytearray(blob)
    ifd = struct.unpack_from("<I", b, 4)[0]
    for i in range(struct.unpack_from("<H", b, ifd)[0]):
        e = ifd + 2 + 12 * i
        if struct.unpack_from("<H", b, e)[0] == tag:
            typ = struct.unpack_from("<H", b, e + 2)[0]
            struct.pack_into("<H" if typ == 3 else "<I", b, e + 8, value)
            return bytes(b)
    raise KeyError(tag)


def test_crafted_dimensions_are_rejected(tmp_path):
    """Width and height come from the file: a frame whose pixel data cannot fit the file, whose
    width * height wraps 32 bits, or whose strips do not cover every row must be refused
    instead of copied past a short allocation."""
    rng = np.random.default_rng(3)
    raster = rng.random((6, 5, 3), dtype=np.float32)
    z = rng.random((6, 5), dtype=np.float32)
    good = str(tmp_path / "good.tif")
    synth_tiff(good, raster, z, "<", 1, True, False, True)
    b = open(good, "rb").read()
    host.load_raw(good)
    cases = {
        "wide.tif": _patch_ifd(b, 256, 0x80000000),                           # 12 * w * h wraps a 64-bit size_t with the next one
        "wrap.tif": _patch_ifd(_patch_ifd(b, 256, 0x80000000), 257, 0x80000000),
        "wrap32.tif": _patch_ifd(_patch_ifd(b, 256, 0x10000), 257, 0x10000),  # w * h == 0 mod 2^32
        "tall.tif": _patch_ifd(b, 257, 7),                                    # one row more than there are strips
    }
    for name, data in cases.items():
        q = str(tmp_path / name)
        open(q, "wb").write(data)
        with pytest.raises(host.SceneError):
            host.load_raw(q)
    # fewer strips than rows: rows_per_strip says 1, the strip table holds 6 entries, height claims 6 but
    # the table is cut to 5 by patching its count
    bb = bytearray(b)
    ifd = struct.unpack_from("<I", bb, 4)[0]
    for i in range(struct.unpack_from("<H", bb, ifd)[0]):
        e = ifd + 2 + 12 * i
        if struct.unpack_from("<H", bb, e)[0] in (273, 279):
            struct.pack_into("<I", bb, e + 4, 5)
    q = str(tmp_path / "fewstrips.tif")
    open(q, "wb").write(bytes(bb))
    with pytest.raises(host.SceneError, match="rows"):
        host.load_raw(q)


def test_postprocess_cli_usage_without_gpu(tmp_path):
    """Argument handling of the `postprocess` executable that ends before any GPU work
    (reference src/postprocess/main.c:46-60)."""
    exe = os.path.join(ROOT, "postprocess")
    if not os.path.exists(exe):
        pytest.skip("postprocess not built (run __graft_entry__.build())")
    r = subprocess.run([exe, "--help"], capture_output=True, text=True)
    assert r.returncode == 0 and "--dof-camera" in r.stdout and "--mist" in r.stdout
    r = subprocess.run([exe, "only-one-argument"], capture_output=True, text=True)
    assert r.returncode == 1 and "Too few arguments" in r.stdout
    r = subprocess.run([exe, str(tmp_path / "missing.tif"), str(tmp_path / "o.tif")], capture_output=True, text=True)
    assert r.returncode == 1 and "Failed to open input file" in r.stdout
    p = str(tmp_path / "eight.tif")
    host.save_tiff(p, np.zeros((4, 4, 3), np.float32), raw=False)
    r = subprocess.run([exe, p, str(tmp_path / "o.tif")], capture_output=True, text=True)
    assert r.returncode == 1 and "32 bits per sample" in r.stdout
