"""Deterministic synthetic scenes for BASELINE.json's configs.  Pure numpy, integer-hash
based, so every process (CPU oracle, each GPU rank) builds byte-identical inputs.  These are INPUTS: the
same arrays are handed to the CUDA backend and to the CPU oracle.

  c1: the 38 spheres of the reference's App::App() (src/App.cpp:101-122)
  c2: Cornell box with a mirror block, a glass sphere and a ceiling area light (~5.3 k triangles)
  c3: 708x708-quad fBm height field, 1 002 528 triangles, GGX-coated material
  c4: 200 000 displaced rocks of 100 triangles each + ground = 20 000 002 triangles, flattened
"""
from __future__ import annotations

import numpy as np

MATERIAL_DTYPE = np.dtype([("albedo", np.float32, 3), ("type", np.uint32), ("emission", np.float32, 3),
                           ("param", np.float32)])
LAMBERT, MIRROR, DIELECTRIC, GGX_COATED = 0, 1, 2, 3


def material(albedo, type=LAMBERT, emission=(0, 0, 0), param=0.0):
    m = np.zeros((), MATERIAL_DTYPE)
    m["albedo"] = albedo
    m["type"] = type
    m["emission"] = emission
    m["param"] = param
    return m


def app_spheres() -> np.ndarray:
    """src/App.cpp:101-122, in creation order (NodeIDs 2..39 in a fresh process)."""
    s = [[0.0, -1.0, 5.0, 1.0], [0.0, -102.0, 5.0, 100.0]]
    for x in range(-5, 6, 2):
        for y in range(-5, 6, 2):
            s.append([float(x), float(y), 10.0, 0.5])
    return np.array(s, np.float32)


def _hash_u32(*keys) -> np.ndarray:
    """Vectorised 32-bit integer hash (xorshift-multiply rounds) of integer arrays."""
    h = np.uint64(0x9E3779B9)
    acc = None
    for k in keys:
        k = np.asarray(k).astype(np.uint64) & np.uint64(0xFFFFFFFF)
        acc = k if acc is None else (acc * np.uint64(0x85EBCA6B) + k) & np.uint64(0xFFFFFFFF)
        acc = (acc ^ (acc >> np.uint64(16))) * np.uint64(0x7FEB352D) & np.uint64(0xFFFFFFFF)
        acc = (acc ^ (acc >> np.uint64(15))) * np.uint64(0x846CA68B) & np.uint64(0xFFFFFFFF)
        acc = acc ^ (acc >> np.uint64(16))
    acc = (acc + h) & np.uint64(0xFFFFFFFF)
    acc = (acc ^ (acc >> np.uint64(16))) * np.uint64(0x7FEB352D) & np.uint64(0xFFFFFFFF)
    acc = acc ^ (acc >> np.uint64(15))
    return acc.astype(np.uint32)


def _hash01(*keys) -> np.ndarray:
    return (_hash_u32(*keys) >> np.uint32(8)).astype(np.float64) * (1.0 / 16777216.0)


def _value_noise(x, z, seed):
    ix = np.floor(x).astype(np.int64)
    iz = np.floor(z).astype(np.int64)
    fx = x - ix
    fz = z - iz
    sx = fx * fx * (3 - 2 * fx)
    sz = fz * fz * (3 - 2 * fz)
    a = _hash01(ix, iz, seed)
    b = _hash01(ix + 1, iz, seed)
    c = _hash01(ix, iz + 1, seed)
    d = _hash01(ix + 1, iz + 1, seed)
    return (a * (1 - sx) + b * sx) * (1 - sz) + (c * (1 - sx) + d * sx) * sz


def _fbm(x, z, seed, octaves=5):
    amp, freq, total, norm = 1.0, 0.35, 0.0, 0.0
    for o in range(octaves):
        total = total + amp * _value_noise(x * freq + 13.7 * o, z * freq - 7.3 * o, seed + o)
        norm += amp
        amp *= 0.5
        freq *= 2.0
    return total / norm


def heightfield(n_quads=708, seed=0x5EED0003, x_range=(-8.0, 8.0), z_range=(2.0, 18.0), base_y=-2.0, amplitude=1.0):
    """c3: (n_quads+1)^2 vertices, 2*n_quads^2 triangles (708 -> 1 002 528)."""
    n = n_quads + 1
    xs = np.linspace(x_range[0], x_range[1], n)
    zs = np.linspace(z_range[0], z_range[1], n)
    X, Z = np.meshgrid(xs, zs, indexing="xy")  # [iz, ix]
    Y = base_y + amplitude * (2.0 * _fbm(X, Z, seed) - 1.0)
    v = np.stack([X, Y, Z], -1).reshape(-1, 3).astype(np.float32)
    iz, ix = np.meshgrid(np.arange(n_quads), np.arange(n_quads), indexing="ij")
    i00 = (iz * n + ix).ravel()
    i10 = i00 + 1
    i01 = i00 + n
    i11 = i01 + 1
    # wound so that cross(e1, e2) points up (+y)
    t0 = np.stack([i00, i01, i10], -1)
    t1 = np.stack([i10, i01, i11], -1)
    idx = np.stack([t0, t1], 1).reshape(-1, 3).astype(np.uint32)
    mats = np.array([material((0.7, 0.7, 0.7), GGX_COATED, param=0.3)], MATERIAL_DTYPE)
    mid = np.zeros(len(idx), np.uint32)
    return v, idx, mid, mats


def _quad(p0, p1, p2, p3):
    """Two triangles (p0,p1,p2), (p0,p2,p3)."""
    return [p0, p1, p2, p3], [[0, 1, 2], [0, 2, 3]]


def _uv_sphere(center, radius, stacks, slices):
    verts = [[0.0, 1.0, 0.0]]
    for i in range(1, stacks):
        th = np.pi * i / stacks
        for j in range(slices):
            ph = 2 * np.pi * j / slices
            verts.append([np.sin(th) * np.cos(ph), np.cos(th), np.sin(th) * np.sin(ph)])
    verts.append([0.0, -1.0, 0.0])
    tris = []
    ring = lambda i, j: 1 + (i - 1) * slices + (j % slices)
    bottom = len(verts) - 1
    for j in range(slices):
        tris.append([0, ring(1, j + 1), ring(1, j)])
        tris.append([bottom, ring(stacks - 1, j), ring(stacks - 1, j + 1)])
    for i in range(1, stacks - 1):
        for j in range(slices):
            a, b, c, d = ring(i, j), ring(i, j + 1), ring(i + 1, j), ring(i + 1, j + 1)
            tris.append([a, b, c])
            tris.append([b, d, c])
    v = np.array(verts) * radius + np.array(center)
    t = np.array(tris)
    # enforce outward winding
    n = np.cross(v[t[:, 1]] - v[t[:, 0]], v[t[:, 2]] - v[t[:, 0]])
    cen = v[t].mean(1) - np.array(center)
    flip = (n * cen).sum(1) < 0
    t[flip] = t[flip][:, [0, 2, 1]]
    return v, t


def cornell_box(sphere_res=50):
    """c2: box x,y in [-1,1], z in [3,5] seen by the reference camera at the origin looking down +Z."""
    mats = np.array([
        material((0.725, 0.71, 0.68)),                       # 0 white
        material((0.63, 0.065, 0.05)),                       # 1 red (left)
        material((0.14, 0.45, 0.091)),                       # 2 green (right)
        material((0.0, 0.0, 0.0), emission=(17.0, 12.0, 4.0)),  # 3 light
        material((0.9, 0.9, 0.9), MIRROR),                   # 4 mirror
        material((1.0, 1.0, 1.0), DIELECTRIC, param=1.5),    # 5 glass
    ], MATERIAL_DTYPE)
    V, T, M = [], [], []

    def add(vs, ts, m):
        base = sum(len(x) for x in V)
        V.append(np.asarray(vs, np.float64))
        T.append(np.asarray(ts, np.int64) + base)
        M.append(np.full(len(ts), m, np.uint32))

    x0, x1, y0, y1, z0, z1 = -1.0, 1.0, -1.0, 1.0, 3.0, 5.0
    add(*_quad([x0, y0, z0], [x1, y0, z0], [x1, y0, z1], [x0, y0, z1]), 0)  # floor
    add(*_quad([x0, y1, z0], [x0, y1, z1], [x1, y1, z1], [x1, y1, z0]), 0)  # ceiling
    add(*_quad([x0, y0, z1], [x1, y0, z1], [x1, y1, z1], [x0, y1, z1]), 0)  # back
    add(*_quad([x0, y0, z0], [x0, y0, z1], [x0, y1, z1], [x0, y1, z0]), 1)  # left
    add(*_quad([x1, y0, z0], [x1, y1, z0], [x1, y1, z1], [x1, y0, z1]), 2)  # right
    # ceiling light, front face (cross(e1,e2)) pointing down
    ly = 0.998
    add(*_quad([-0.25, ly, 3.75], [0.25, ly, 3.75], [0.25, ly, 4.25], [-0.25, ly, 4.25]), 3)
    # mirror block: rotated box
    c, s = np.cos(np.radians(-18.0)), np.sin(np.radians(-18.0))
    hx, hy, hz = 0.3, 0.6, 0.3
    cx, cy, cz = -0.38, -0.4, 4.35
    corners = []
    for dy in (-hy, hy):
        for dx, dz in ((-hx, -hz), (hx, -hz), (hx, hz), (-hx, hz)):
            corners.append([cx + c * dx + s * dz, cy + dy, cz - s * dx + c * dz])
    b = corners
    faces = [[0, 1, 2, 3], [4, 7, 6, 5], [0, 4, 5, 1], [1, 5, 6, 2], [2, 6, 7, 3], [3, 7, 4, 0]]
    for f in faces:
        add(*_quad(b[f[0]], b[f[1]], b[f[2]], b[f[3]]), 4)
    sv, st = _uv_sphere((0.45, -0.65, 3.75), 0.35, sphere_res, sphere_res)
    add(sv, st, 5)
    v = np.concatenate(V).astype(np.float32)
    t = np.concatenate(T).astype(np.uint32)
    m = np.concatenate(M)
    return v, t, m, mats


def _icosphere(level=1):
    p = (1 + 5 ** 0.5) / 2
    v = [[-1, p, 0], [1, p, 0], [-1, -p, 0], [1, -p, 0], [0, -1, p], [0, 1, p], [0, -1, -p], [0, 1, -p],
         [p, 0, -1], [p, 0, 1], [-p, 0, -1], [-p, 0, 1]]
    f = [[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6],
         [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10],
         [8, 6, 7], [9, 8, 1]]
    v = [np.array(x, np.float64) / np.linalg.norm(x) for x in v]
    for _ in range(level):
        cache, nf = {}, []

        def mid(a, b):
            key = (min(a, b), max(a, b))
            if key not in cache:
                m = v[a] + v[b]
                v.append(m / np.linalg.norm(m))
                cache[key] = len(v) - 1
            return cache[key]

        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [[a, ab, ca], [b, bc, ab], [c, ca, bc], [ab, bc, ca]]
        f = nf
    return np.array(v), np.array(f)


def rock_field(n_instances=200_000, seed=0xC4, extent_x=(-50.0, 50.0), extent_z=(2.0, 102.0), ground_y=-2.0):
    """c4: n_instances displaced rocks (80-triangle icosphere + 20-triangle skirt) on a jittered grid,
    flattened to world-space triangles, plus a 2-triangle ground quad."""
    iv, it = _icosphere(1)                     # 42 verts, 80 tris
    # 20-triangle skirt: a 10-gon ring at the base, two triangles per side down to a wider ring
    k = 10
    ang = 2 * np.pi * np.arange(k) / k
    r0 = np.stack([0.9 * np.cos(ang), np.full(k, -0.55), 0.9 * np.sin(ang)], -1)
    r1 = np.stack([1.35 * np.cos(ang), np.full(k, -1.0), 1.35 * np.sin(ang)], -1)
    bv = np.concatenate([iv, r0, r1])
    o0, o1 = len(iv), len(iv) + k
    skirt = []
    for j in range(k):
        a, b = o0 + j, o0 + (j + 1) % k
        c, d = o1 + j, o1 + (j + 1) % k
        skirt += [[a, b, c], [b, d, c]]
    bt = np.concatenate([it, np.array(skirt)])
    nvb, ntb = len(bv), len(bt)
    assert ntb == 100
    gx = int(np.ceil(np.sqrt(n_instances)))
    ids = np.arange(n_instances)
    cellx = (extent_x[1] - extent_x[0]) / gx
    cellz = (extent_z[1] - extent_z[0]) / gx
    px = extent_x[0] + (ids % gx + 0.15 + 0.7 * _hash01(ids, 1, seed)) * cellx
    pz = extent_z[0] + (ids // gx + 0.15 + 0.7 * _hash01(ids, 2, seed)) * cellz
    scale = (0.25 + 0.2 * _hash01(ids, 3, seed)) * min(cellx, cellz)
    rot = 2 * np.pi * _hash01(ids, 4, seed)
    cr, sr = np.cos(rot), np.sin(rot)
    # per-instance, per-vertex radial displacement
    disp = 1.0 + 0.35 * (_hash01(ids[:, None], np.arange(nvb)[None, :], seed + 5) - 0.5)
    disp[:, len(iv):] = 1.0
    local = bv[None, :, :] * disp[:, :, None]
    wx = (cr[:, None] * local[:, :, 0] + sr[:, None] * local[:, :, 2]) * scale[:, None] + px[:, None]
    wz = (-sr[:, None] * local[:, :, 0] + cr[:, None] * local[:, :, 2]) * scale[:, None] + pz[:, None]
    wy = (local[:, :, 1] + 1.0) * scale[:, None] + ground_y
    v = np.stack([wx, wy, wz], -1).reshape(-1, 3).astype(np.float32)
    t = (bt[None, :, :] + (ids * nvb)[:, None, None]).reshape(-1, 3).astype(np.uint32)
    mid = (_hash_u32(ids, 9, seed) % np.uint32(3)).astype(np.uint32).repeat(ntb) + np.uint32(1)
    # ground quad (material 0)
    g0 = len(v)
    gv = np.array([[extent_x[0] - 10, ground_y, extent_z[0] - 10], [extent_x[0] - 10, ground_y, extent_z[1] + 10],
                   [extent_x[1] + 10, ground_y, extent_z[1] + 10], [extent_x[1] + 10, ground_y, extent_z[0] - 10]],
                  np.float32)
    v = np.concatenate([v, gv])
    t = np.concatenate([t, np.array([[g0, g0 + 1, g0 + 2], [g0, g0 + 2, g0 + 3]], np.uint32)])
    mid = np.concatenate([mid, np.zeros(2, np.uint32)])
    mats = np.array([
        material((0.55, 0.5, 0.45), GGX_COATED, param=0.5),
        material((0.7, 0.7, 0.7), GGX_COATED, param=0.3),
        material((0.6, 0.35, 0.25), GGX_COATED, param=0.2),
        material((0.35, 0.4, 0.6), GGX_COATED, param=0.4),
    ], MATERIAL_DTYPE)
    return v, t, mid, mats


def octahedral_environment(size=64, sun_dir=(0.4, 0.7, 0.6), sun_power=40.0, seed=0xE0):
    """Synthetic HDR environment as a size x size octahedral map (pole +Y, layout of cudapt_upload_environment):
    the reference's sky gradient (CPT.cpp:286-292) plus a bright sun lobe and a little hashed texture, so that a
    wrong texel address or weight shows up in parity tests."""
    j, i = np.meshgrid(np.arange(size), np.arange(size), indexing="ij")
    u = (i + 0.5) / size * 2.0 - 1.0  # p.x
    v = (j + 0.5) / size * 2.0 - 1.0  # p.z
    y = 1.0 - np.abs(u) - np.abs(v)
    x, z = u.copy(), v.copy()
    lower = y < 0
    x[lower] = (1.0 - np.abs(v[lower])) * np.where(u[lower] < 0, -1.0, 1.0)
    z[lower] = (1.0 - np.abs(u[lower])) * np.where(v[lower] < 0, -1.0, 1.0)
    d = np.stack([x, y, z], -1)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    t = 0.5 * (d[..., 1] + 1.0)
    sky = (1.0 - t)[..., None] * np.array([1.0, 1.0, 1.0]) + t[..., None] * np.array([0.5, 0.7, 1.0])
    s = np.asarray(sun_dir, np.float64)
    s /= np.linalg.norm(s)
    lobe = np.clip(d @ s, 0.0, 1.0) ** 64
    sun = sun_power * lobe[..., None] * np.array([1.0, 0.9, 0.7])
    grain = 0.1 * _hash01(i.astype(np.uint32), j.astype(np.uint32), np.uint32(seed))[..., None]
    return np.ascontiguousarray(sky + sun + grain, np.float32)
