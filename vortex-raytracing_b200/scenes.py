"""Procedural scenes for the BASELINE.json configs — synthetic input generation only.

Every generator is a pure function of its arguments and of the reference's integer hash
(WangHash, tests/regression/raycast/common.h:85-91), evaluated in numpy uint32/float32 so the same bytes come
out on every machine.  A mesh is a dict:
    tris (n,9) f32  : v0 v1 v2                         (tri_t,    geometry.h:1401-1405)
    triex (n,15) f32: N0 N1 N2 uv0 uv1 uv2             (tri_ex_t, raycast/common.h:19-22)
    tex (h,w) u32   : 0x00RRGGBB texels                (surface.cpp:47-50)
    reflectivity    : float                            (blas_node_t.reflectivity)
"""
import numpy as np


def wang_hash(s):
    s = np.asarray(s, dtype=np.uint32)
    with np.errstate(over="ignore"):
        s = (s ^ np.uint32(61)) ^ (s >> np.uint32(16))
        s = s * np.uint32(9)
        s = s ^ (s >> np.uint32(4))
        s = s * np.uint32(0x27D4EB2D)
        s = s ^ (s >> np.uint32(15))
    return s


def _lattice(ix, iz, octave, seed):
    with np.errstate(over="ignore"):
        k = (ix.astype(np.uint32) * np.uint32(73856093)) ^ (iz.astype(np.uint32) * np.uint32(19349663)) ^ np.uint32(
            (octave * 83492791 + seed * 2654435761) & 0xFFFFFFFF)
    return (wang_hash(k) >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / (1 << 24))


def value_noise(x, z, octaves=3, seed=1, base_freq=4.0):
    """Sum of `octaves` octaves of hashed value noise over the plane; float32 throughout."""
    out = np.zeros_like(x, dtype=np.float32)
    amp, freq = np.float32(1.0), np.float32(base_freq)
    for o in range(octaves):
        fx, fz = x * freq, z * freq
        x0, z0 = np.floor(fx), np.floor(fz)
        tx, tz = (fx - x0).astype(np.float32), (fz - z0).astype(np.float32)
        sx, sz = tx * tx * (np.float32(3) - np.float32(2) * tx), tz * tz * (np.float32(3) - np.float32(2) * tz)
        ix, iz = x0.astype(np.int64) & 0xFFFFFFFF, z0.astype(np.int64) & 0xFFFFFFFF
        a, b = _lattice(ix, iz, o, seed), _lattice(ix + 1, iz, o, seed)
        c, d = _lattice(ix, iz + 1, o, seed), _lattice(ix + 1, iz + 1, o, seed)
        top = a + (b - a) * sx
        bot = c + (d - c) * sx
        out = out + amp * (top + (bot - top) * sz)
        amp, freq = amp * np.float32(0.5), freq * np.float32(2)
    return out.astype(np.float32)


def noise_texture(size=256, seed=1):
    """size x size 0x00RRGGBB texture: earthy value-noise colours with a faint grid."""
    u = (np.arange(size, dtype=np.float32) + np.float32(0.5)) / np.float32(size)
    uu, vv = np.meshgrid(u, u)
    n = value_noise(uu, vv, octaves=3, seed=seed + 17, base_freq=8.0) / np.float32(1.75)
    r = np.clip(80 + 150 * n, 0, 255).astype(np.uint32)
    g = np.clip(110 + 120 * (1 - n), 0, 255).astype(np.uint32)
    b = np.clip(60 + 90 * n, 0, 255).astype(np.uint32)
    grid = ((np.arange(size) % 32) == 0)
    r[grid, :] //= 2
    r[:, grid] //= 2
    return ((r << 16) | (g << 8) | b).astype(np.uint32)


def heightfield(n, height=0.15, seed=1, tex_size=256, reflectivity=0.0):
    """n x n quads (2 n^2 triangles) over [-1,1]^2 in XZ, height along +Y = height * 3 octaves of value noise.
    224 -> 100 352 tris, 708 -> 1 002 528, 2237 -> 10 008 338."""
    g = np.linspace(-1.0, 1.0, n + 1, dtype=np.float32)
    X, Z = np.meshgrid(g, g, indexing="xy")            # [iz, ix]
    Y = (np.float32(height) * value_noise(X, Z, 3, seed)).astype(np.float32)
    # vertex normals from central differences of the height grid
    dydx = np.zeros_like(Y)
    dydz = np.zeros_like(Y)
    step = np.float32(2.0 / n)
    dydx[:, 1:-1] = (Y[:, 2:] - Y[:, :-2]) / (2 * step)
    dydx[:, 0] = (Y[:, 1] - Y[:, 0]) / step
    dydx[:, -1] = (Y[:, -1] - Y[:, -2]) / step
    dydz[1:-1, :] = (Y[2:, :] - Y[:-2, :]) / (2 * step)
    dydz[0, :] = (Y[1, :] - Y[0, :]) / step
    dydz[-1, :] = (Y[-1, :] - Y[-2, :]) / step
    N = np.stack([-dydx, np.ones_like(Y), -dydz], axis=-1).astype(np.float32)
    N /= np.sqrt((N * N).sum(-1, keepdims=True)).astype(np.float32)
    P = np.stack([X, Y, Z], axis=-1).astype(np.float32)
    UV = np.stack([(X + 1) * np.float32(0.5), (Z + 1) * np.float32(0.5)], axis=-1).astype(np.float32)

    def quad(a):   # corners of every quad: 00 10 01 11 (ix, iz offsets)
        return a[:-1, :-1], a[:-1, 1:], a[1:, :-1], a[1:, 1:]

    p00, p10, p01, p11 = quad(P)
    n00, n10, n01, n11 = quad(N)
    t00, t10, t01, t11 = quad(UV)
    # two triangles per quad, counter-clockwise seen from +Y
    tris = np.stack([np.concatenate([p00, p01, p11], -1), np.concatenate([p00, p11, p10], -1)], axis=2)
    nrm = np.stack([np.concatenate([n00, n01, n11], -1), np.concatenate([n00, n11, n10], -1)], axis=2)
    uvs = np.stack([np.concatenate([t00, t01, t11], -1), np.concatenate([t00, t11, t10], -1)], axis=2)
    tris = np.ascontiguousarray(tris.reshape(-1, 9), dtype=np.float32)
    triex = np.ascontiguousarray(np.concatenate([nrm.reshape(-1, 9), uvs.reshape(-1, 6)], axis=1), dtype=np.float32)
    return {"tris": tris, "triex": triex, "tex": noise_texture(tex_size, seed), "reflectivity": float(reflectivity)}


def _quad(p0, p1, p2, p3):
    """Two triangles (p0 p1 p2) (p0 p2 p3) with the flat normal and unit-square UVs."""
    p0, p1, p2, p3 = [np.asarray(p, dtype=np.float32) for p in (p0, p1, p2, p3)]
    n = np.cross(p1 - p0, p2 - p0).astype(np.float32)
    n = (n / np.float32(np.sqrt((n * n).sum()))).astype(np.float32)
    uv = [np.array(u, dtype=np.float32) for u in ((0, 0), (1, 0), (1, 1), (0, 1))]
    tris = [np.concatenate([p0, p1, p2]), np.concatenate([p0, p2, p3])]
    ex = [np.concatenate([n, n, n, uv[0], uv[1], uv[2]]), np.concatenate([n, n, n, uv[0], uv[2], uv[3]])]
    return tris, ex


def _box(lo, hi, yaw=0.0):
    """5 visible faces (no bottom) of an axis-aligned box rotated by `yaw` about its vertical axis: 10 triangles."""
    lo, hi = np.asarray(lo, np.float32), np.asarray(hi, np.float32)
    c = (lo + hi) * np.float32(0.5)
    cs, sn = np.float32(np.cos(yaw)), np.float32(np.sin(yaw))

    def rot(p):
        q = np.asarray(p, np.float32) - c
        return np.array([q[0] * cs + q[2] * sn + c[0], p[1], -q[0] * sn + q[2] * cs + c[2]], dtype=np.float32)

    x0, y0, z0 = lo
    x1, y1, z1 = hi
    faces = [((x0, y1, z0), (x0, y1, z1), (x1, y1, z1), (x1, y1, z0)),   # top
             ((x0, y0, z0), (x0, y1, z0), (x1, y1, z0), (x1, y0, z0)),   # front (-z)
             ((x1, y0, z1), (x1, y1, z1), (x0, y1, z1), (x0, y0, z1)),   # back  (+z)
             ((x0, y0, z1), (x0, y1, z1), (x0, y1, z0), (x0, y0, z0)),   # left  (-x)
             ((x1, y0, z0), (x1, y1, z0), (x1, y1, z1), (x1, y0, z1))]   # right (+x)
    T, E = [], []
    for f in faces:
        t, e = _quad(*[rot(p) for p in f])
        T += t
        E += e
    return T, E


def cornell_box(reflectivity=0.0):
    """32 triangles in [-1,1]^3: floor, ceiling, back, left, right walls (10) + ceiling light quad (2) + short box (10)
    + tall box (10).  The camera looks down +Z (Scene::computeFramingCamera puts it at -Z), so the open side is -Z.
    Flat walls give zero-extent boxes — the quantisation edge case of the RT-unit path."""
    T, E = [], []
    walls = [((-1, -1, -1), (1, -1, -1), (1, -1, 1), (-1, -1, 1)),      # floor  (normal +y)
             ((-1, 1, 1), (1, 1, 1), (1, 1, -1), (-1, 1, -1)),          # ceiling (normal -y)
             ((-1, -1, 1), (1, -1, 1), (1, 1, 1), (-1, 1, 1)),          # back wall (normal -z)
             ((-1, -1, -1), (-1, -1, 1), (-1, 1, 1), (-1, 1, -1)),      # left  (normal +x)
             ((1, -1, 1), (1, -1, -1), (1, 1, -1), (1, 1, 1)),          # right (normal -x)
             ((-0.25, 0.999, 0.25), (0.25, 0.999, 0.25), (0.25, 0.999, -0.25), (-0.25, 0.999, -0.25))]   # light
    for w in walls:
        t, e = _quad(*w)
        T += t
        E += e
    for lo, hi, yaw in (((-0.65, -1.0, -0.45), (-0.05, -0.4, 0.15), 0.3), ((0.1, -1.0, 0.05), (0.7, 0.2, 0.65), -0.3)):
        t, e = _box(lo, hi, yaw)
        T += t
        E += e
    tris = np.ascontiguousarray(np.stack(T), dtype=np.float32)
    triex = np.ascontiguousarray(np.stack(E), dtype=np.float32)
    assert tris.shape == (32, 9)
    # 4x4 texture: warm white with red/green accents so walls differ
    tex = np.full((4, 4), 0x00C8C0B0, dtype=np.uint32)
    tex[0, :] = 0x00C04030
    tex[3, :] = 0x0030A040
    return {"tris": tris, "triex": triex, "tex": tex, "reflectivity": float(reflectivity)}


def icosphere(center, radius, subdiv=2, tex=None, reflectivity=0.0):
    """Subdivided icosahedron (20 * 4^subdiv triangles) with smooth normals and spherical UVs."""
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = np.array([(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t),
                  (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)], dtype=np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array([(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6),
                  (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10),
                  (8, 6, 7), (9, 8, 1)])
    tri = v[f]                                             # (20,3,3)
    for _ in range(subdiv):
        a, b, c = tri[:, 0], tri[:, 1], tri[:, 2]
        ab, bc, ca = a + b, b + c, c + a
        ab /= np.linalg.norm(ab, axis=1, keepdims=True)
        bc /= np.linalg.norm(bc, axis=1, keepdims=True)
        ca /= np.linalg.norm(ca, axis=1, keepdims=True)
        tri = np.concatenate([np.stack([a, ab, ca], 1), np.stack([b, bc, ab], 1), np.stack([c, ca, bc], 1), np.stack([ab, bc, ca], 1)])
    n = tri.astype(np.float32)
    p = (np.asarray(center, np.float32) + np.float32(radius) * n).astype(np.float32)
    uv = np.stack([np.arctan2(n[..., 2], n[..., 0]) / (2 * np.pi) + 0.5, np.arccos(np.clip(n[..., 1], -1, 1)) / np.pi], -1).astype(np.float32)
    tris = np.ascontiguousarray(p.reshape(-1, 9))
    triex = np.ascontiguousarray(np.concatenate([n.reshape(-1, 9), uv.reshape(-1, 6)], axis=1).astype(np.float32))
    if tex is None:
        tex = noise_texture(64, 7)
    return {"tris": tris, "triex": triex, "tex": tex, "reflectivity": float(reflectivity)}


def random_rays(n, bmin, bmax, seed=1, start=0):
    """Incoherent rays (config 5): origin uniform in the box, direction uniform on the sphere
    (z = 1 - 2 u1, phi = 2 pi u2), all from WangHash(i)-seeded integers; ray indices start .. start+n-1 (so that ranks
    can generate their own shard of one global ray set).  Returns (n,6) float32."""
    i = np.arange(start, start + n, dtype=np.uint64).astype(np.uint32)
    inv = np.float32(1.0 / (1 << 24))

    def u(k):
        with np.errstate(over="ignore"):
            return (wang_hash(i * np.uint32(6) + np.uint32(k) + np.uint32(seed * 977)) >> np.uint32(8)).astype(np.float32) * inv

    bmin, bmax = np.asarray(bmin, np.float32), np.asarray(bmax, np.float32)
    o = np.stack([bmin[a] + (bmax[a] - bmin[a]) * u(a) for a in range(3)], axis=1)
    z = np.float32(1) - np.float32(2) * u(3)
    phi = np.float32(2 * np.pi) * u(4)
    r = np.sqrt(np.maximum(np.float32(0), np.float32(1) - z * z)).astype(np.float32)
    d = np.stack([r * np.cos(phi), r * np.sin(phi), z], axis=1).astype(np.float32)
    return np.ascontiguousarray(np.concatenate([o, d], axis=1), dtype=np.float32)
