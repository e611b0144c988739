#!/usr/bin/env python3
"""Deterministic stand-in meshes for the two STL files that are missing from the
reference checkout (/root/reference/.MISSING_LARGE_BLOBS: meshes/dragon.stl,
meshes/menger_sponge.stl).  Both are written as binary STL (80-byte
header that does not start with "solid", u32 count, 50-byte records) at the paths
scenes/scene5.json and scenes/scene6.json name, so the CPU reference and the GPU
build read the same file.

  dragon.stl         closed tube around a trefoil knot with a tapering, scaled
                     ("bumpy") cross-section: nu x nv quads -> 2 nu nv triangles.
                     929 x 469 gives 871 402 triangles (the Stanford dragon it stands in for has 871 414).
                     NOT the Stanford geometry.
  menger_sponge.stl  level-n Menger sponge in [-0.5, 0.5]^3, exposed faces only
                     (watertight, no coincident interior faces):
                     2 * (2 * 20^n + 4 * 8^n) triangles -> level 4: 672 768.

Small variants (dragon_small.stl, menger_small.stl) are used by the tests.
Usage: tools/gen_meshes.py [--force] [--only small|large]
"""
import hashlib
import os
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def write_stl(path, tris, title):
    """tris: (n, 3, 3) float32 vertices, outward winding."""
    tris = np.ascontiguousarray(tris, dtype=np.float32)
    n = len(tris)
    e0 = tris[:, 1] - tris[:, 0]
    e1 = tris[:, 2] - tris[:, 0]
    nrm = np.cross(e0, e1)
    ln = np.linalg.norm(nrm, axis=1, keepdims=True)
    nrm = (nrm / np.maximum(ln, 1e-30)).astype(np.float32)
    rec = np.zeros(n, dtype=np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("a", "<u2")]))
    rec["n"] = nrm
    rec["v"] = tris
    header = (b"binary stl; " + title.encode())[:80].ljust(80, b" ")
    assert not header.startswith(b"solid")
    with open(path, "wb") as f:
        f.write(header)
        f.write(struct.pack("<I", n))
        f.write(rec.tobytes())
    return n


def menger_cells(level):
    n = 3 ** level
    idx = np.arange(n)
    filled = np.ones((n, n, n), dtype=bool)
    x, y, z = np.meshgrid(idx, idx, idx, indexing="ij")
    for _ in range(level):
        ones = (x % 3 == 1).astype(np.int8) + (y % 3 == 1).astype(np.int8) + (z % 3 == 1).astype(np.int8)
        filled &= ones < 2
        x, y, z = x // 3, y // 3, z // 3
    return filled


def menger_sponge(level):
    filled = menger_cells(level)
    n = filled.shape[0]
    h = 1.0 / n
    padded = np.zeros((n + 2,) * 3, dtype=bool)
    padded[1:-1, 1:-1, 1:-1] = filled
    tris = []
    # for each axis and direction: faces of filled cells whose neighbour is empty
    for axis in range(3):
        for sign in (-1, 1):
            sl = [slice(1, -1)] * 3
            sl[axis] = slice(1 + sign, n + 1 + sign)
            exposed = filled & ~padded[tuple(sl)]
            cells = np.argwhere(exposed).astype(np.float64)
            if not len(cells):
                continue
            u, v = (axis + 1) % 3, (axis + 2) % 3
            base = cells * h - 0.5
            if sign > 0:
                base[:, axis] += h
            du = np.zeros(3)
            dv = np.zeros(3)
            du[u] = h
            dv[v] = h
            p00, p10, p11, p01 = base, base + du, base + du + dv, base + dv
            # du x dv = +axis: counter-clockwise seen from +axis
            if sign > 0:
                quads = (p00, p10, p11, p01)
            else:
                quads = (p00, p01, p11, p10)
            a, b, c, d = quads
            tris.append(np.stack([a, b, c], axis=1))
            tris.append(np.stack([a, c, d], axis=1))
    return np.concatenate(tris, axis=0).astype(np.float32)


def rotation_minimising_frames(points, tangents):
    """Double-reflection method (Wang et al. 2008) along a closed polyline; the
    residual twist at the seam is spread uniformly."""
    n = len(points)
    r = np.zeros_like(points)
    t0 = tangents[0]
    ref = np.array([0.0, 0.0, 1.0]) if abs(t0[2]) < 0.9 else np.array([1.0, 0.0, 0.0])
    r[0] = np.cross(t0, ref)
    r[0] /= np.linalg.norm(r[0])
    for i in range(n):
        j = (i + 1) % n
        v1 = points[j] - points[i]
        c1 = v1 @ v1
        rl = r[i] - (2.0 / c1) * (v1 @ r[i]) * v1
        tl = tangents[i] - (2.0 / c1) * (v1 @ tangents[i]) * v1
        v2 = tangents[j] - tl
        c2 = v2 @ v2
        nxt = rl - (2.0 / c2) * (v2 @ rl) * v2 if c2 > 1e-30 else rl
        if j == 0:
            last = nxt
        else:
            r[j] = nxt
    s0 = np.cross(tangents[0], r[0])
    twist = np.arctan2(last @ s0, last @ r[0])
    s = np.cross(tangents, r)
    ang = -twist * np.arange(n) / n
    rr = r * np.cos(ang)[:, None] + s * np.sin(ang)[:, None]
    ss = np.cross(tangents, rr)
    return rr, ss


def dragon_standin(nu, nv):
    t = np.linspace(0.0, 2.0 * np.pi, nu, endpoint=False)
    # trefoil knot, model z is "up" (scene5 rotates it onto -y)
    cx = 7.5 * (np.sin(t) + 2.0 * np.sin(2.0 * t))
    cy = 6.5 * np.sin(3.0 * t)
    cz = 5.0 * (np.cos(t) - 2.0 * np.cos(2.0 * t)) + 1.5
    c = np.stack([cx, cy, cz], axis=1)
    d = np.stack([7.5 * (np.cos(t) + 4.0 * np.cos(2.0 * t)), 19.5 * np.cos(3.0 * t),
                  5.0 * (-np.sin(t) + 4.0 * np.sin(2.0 * t))], axis=1)
    tan = d / np.linalg.norm(d, axis=1, keepdims=True)
    nrm, bin_ = rotation_minimising_frames(c, tan)
    v = np.linspace(0.0, 2.0 * np.pi, nv, endpoint=False)
    # body radius: thick "torso", thin "tail", periodic in t
    body = 2.3 + 1.1 * np.cos(t - 0.6) + 0.35 * np.cos(2.0 * t + 1.0)
    tt, vv = np.meshgrid(t, v, indexing="ij")
    scales = 1.0 + 0.07 * np.sin(46.0 * tt) * np.sin(14.0 * vv) + 0.04 * np.cos(9.0 * tt + 3.0 * vv)
    ridge = 1.0 + 0.22 * np.exp(-((np.mod(vv - np.pi / 2 + np.pi, 2 * np.pi) - np.pi) ** 2) / 0.05)  # dorsal crest
    r = body[:, None] * scales * ridge
    p = c[:, None, :] + r[..., None] * (np.cos(vv)[..., None] * nrm[:, None, :] + np.sin(vv)[..., None] * bin_[:, None, :])
    p = p.astype(np.float32)
    i0 = np.arange(nu)[:, None]
    j0 = np.arange(nv)[None, :]
    i1 = (i0 + 1) % nu
    j1 = (j0 + 1) % nv
    a, b, cc, dd = p[i0, j0], p[i1, j0], p[i1, j1], p[i0, j1]
    t1 = np.stack([a, b, cc], axis=2).reshape(-1, 3, 3)
    t2 = np.stack([a, cc, dd], axis=2).reshape(-1, 3, 3)
    tris = np.concatenate([t1, t2], axis=0)
    # make the winding outward: compare one triangle normal with the radial direction
    n0 = np.cross(tris[0, 1] - tris[0, 0], tris[0, 2] - tris[0, 0])
    radial = tris[0].mean(axis=0) - c[0]
    if n0 @ radial < 0:
        tris = tris[:, ::-1, :]
    return np.ascontiguousarray(tris, dtype=np.float32)


TARGETS = {
    # name: (kind, builder, description)
    "meshes/dragon.stl": ("large", lambda: dragon_standin(929, 469), "trefoil-tube dragon stand-in 929x469"),
    "meshes/menger_sponge.stl": ("large", lambda: menger_sponge(4), "menger sponge level 4 exposed faces"),
    "meshes/dragon_small.stl": ("small", lambda: dragon_standin(160, 64), "trefoil-tube dragon stand-in 160x64"),
    "meshes/menger_small.stl": ("small", lambda: menger_sponge(2), "menger sponge level 2 exposed faces"),
}


def sha256(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 20), b""):
            h.update(blk)
    return h.hexdigest()


def main(argv):
    force = "--force" in argv
    only = argv[argv.index("--only") + 1] if "--only" in argv else None
    for rel, (kind, build, title) in TARGETS.items():
        if only and kind != only:
            continue
        path = os.path.join(ROOT, rel)
        if os.path.exists(path) and not force:
            continue
        os.makedirs(os.path.dirname(path), exist_ok=True)
        n = write_stl(path, build(), title)
        print(f"{rel}: {n} triangles, sha256 {sha256(path)}")
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
