"""Deterministic synthetic voxel worlds: integer-hash noise only, no file I/O.

  terrain(N)      heightfield shell h(x,z) = N/4 + N/8*fbm(x/N, z/N; 5 octaves value noise, seed 0xB10C),
                  filled voxels {h-2 <= y <= h}, 8 materials by height band        (configs 3-5)
  blob_mesh(N)    a solid "voxelised mesh": union of hash-placed ellipsoids, surface shell + sparse interior
                  struts, 16 materials                                             (configs 1-2)
Both return (xyz int32 (n,3), material ids uint32 (n,), materials MAT_DT table); feed them to
WorldSvoGpu.from_voxels to get the reference's upload arrays.
"""
import numpy as np

from .api import MAT_DT


def _hash2(ix, iz, seed):
    h = (ix.astype(np.uint32) * np.uint32(0x9E3779B1)) ^ (iz.astype(np.uint32) * np.uint32(0x85EBCA77)) ^ np.uint32(seed)
    h ^= h >> np.uint32(15)
    h *= np.uint32(0x2C1B3C6D)
    h ^= h >> np.uint32(12)
    h *= np.uint32(0x297A2D39)
    h ^= h >> np.uint32(15)
    return h.astype(np.float64) / 4294967296.0


def _value_noise(x, z, seed):
    ix, iz = np.floor(x).astype(np.int64), np.floor(z).astype(np.int64)
    fx, fz = x - ix, z - iz
    sx, sz = fx * fx * (3 - 2 * fx), fz * fz * (3 - 2 * fz)
    a, b = _hash2(ix, iz, seed), _hash2(ix + 1, iz, seed)
    c, d = _hash2(ix, iz + 1, seed), _hash2(ix + 1, iz + 1, seed)
    return (a * (1 - sx) + b * sx) * (1 - sz) + (c * (1 - sx) + d * sx) * sz


def fbm(u, v, octaves=5, seed=0xB10C):
    s, amp, freq, norm = 0.0, 0.5, 4.0, 0.0
    for o in range(octaves):
        s = s + amp * _value_noise(u * freq, v * freq, seed + 101 * o)
        norm += amp
        amp *= 0.5
        freq *= 2.0
    return s / norm  # in [0,1)


def default_materials(n=16):
    """n+1 MaterialGpu records: id 0 = the reference's default grey (material.cpp:19-29), then a fixed palette;
    one metallic, one dim emissive and one rough entry so every shading branch of raygen.rgen is exercised."""
    m = np.zeros(n + 1, MAT_DT)
    def flags(metal, rough, typ=0):
        return (int(metal * 255) << 24) | (int(rough * 255) << 16) | (typ << 12) | (15 << 8) | int(0.5 * 255)
    m[0]["albedo"] = (0.8, 0.8, 0.8)
    m[0]["flags"] = flags(0.0, 0.5)
    pal = [(0.35, 0.55, 0.25), (0.45, 0.40, 0.30), (0.55, 0.52, 0.50), (0.85, 0.85, 0.90), (0.25, 0.35, 0.65), (0.75, 0.65, 0.40),
           (0.60, 0.25, 0.20), (0.30, 0.30, 0.32), (0.90, 0.75, 0.30), (0.20, 0.60, 0.55), (0.70, 0.40, 0.65), (0.95, 0.95, 0.95),
           (0.15, 0.15, 0.18), (0.50, 0.70, 0.30), (0.80, 0.50, 0.25), (0.40, 0.45, 0.75)]
    for i in range(1, n + 1):
        m[i]["albedo"] = pal[(i - 1) % len(pal)]
        m[i]["flags"] = flags(0.0, 0.6 + 0.05 * ((i * 7) % 7))
    if n >= 8:
        m[8]["flags"] = flags(1.0, 0.2, 1)              # metal
        m[4]["flags"] = flags(0.0, 0.1)                 # glossy dielectric
    if n >= 12:
        m[12]["emission"] = (0.6, 0.5, 0.3)             # dim emissive (continues the path, raygen.rgen:265-277)
        m[12]["ior"] = 1.0
        m[12]["flags"] = flags(0.0, 0.5, 3)
    return m


def terrain_heights(N=2048, seed=0xB10C, rows=512):
    """h[z, x] of the heightfield, int32 (N, N); evaluated in bands of `rows` z-rows so N = 8192 stays within a few GB."""
    xs = np.arange(N, dtype=np.float64)
    h = np.empty((N, N), np.int32)

    def band(z0):
        X, Z = np.meshgrid(xs, xs[z0:z0 + rows], indexing="xy")
        h[z0:z0 + rows] = np.floor(N / 4 + N / 8 * fbm(X / N, Z / N, 5, seed)).astype(np.int32)

    starts = list(range(0, N, rows))
    if len(starts) > 1:  # numpy releases the GIL inside its loops: bands in parallel (elementwise, so identical values)
        from concurrent.futures import ThreadPoolExecutor
        import os
        with ThreadPoolExecutor(min(len(starts), os.cpu_count() or 1, 16)) as ex:
            list(ex.map(band, starts))
    else:
        band(0)
    return h


def terrain(N=2048, seed=0xB10C, thickness=3, origin=(0, 0, 0)):
    """3-voxel heightfield shell over an N x N footprint: filled voxels {h - 2 <= y <= h}."""
    h = terrain_heights(N, seed).ravel()
    n = h.size
    X = np.tile(np.arange(N, dtype=np.int32), N)
    Z = np.repeat(np.arange(N, dtype=np.int32), N)
    xyz = np.empty((thickness, n, 3), np.int32)
    for k in range(thickness):
        xyz[k, :, 0] = X + origin[0]
        xyz[k, :, 1] = h - k + origin[1]
        xyz[k, :, 2] = Z + origin[2]
    xyz = xyz.reshape(-1, 3)
    band = ((h - N // 4) * 8 // max(N // 8, 1)).clip(0, 7).astype(np.uint32) + 1
    return xyz, np.tile(band, thickness), default_materials(16)


def dense_mesh(N=512):
    """The c2 scene: 40 large ellipsoids spread over the whole N^3 box — a solid that fills the frame from the c2 camera."""
    return blob_mesh(N, blobs=40, centres=(0.15, 0.85), radii=(0.10, 0.22))


def blob_mesh(N=512, seed=0x51DE, blobs=24, centres=(0.2, 0.8), radii=(0.06, 0.18)):
    """Solid union of ellipsoids in an N^3 box, voxelised as a 2-voxel surface shell plus every 8th interior
    lattice line (keeps dense regions dense without billions of voxels). centres / radii: ranges, in units of N, the
    ellipsoid centres and semi-axes are drawn from. Each ellipsoid is evaluated inside its own bounding box only
    (outside it the implicit function is >= 1 and can change nothing), z-slabs in parallel."""
    rng = np.random.Generator(np.random.PCG64(seed))
    c = rng.uniform(centres[0] * N, centres[1] * N, (blobs, 3)).astype(np.float32)
    r = rng.uniform(radii[0] * N, radii[1] * N, (blobs, 3)).astype(np.float32)
    lo = np.clip(np.floor(c - r).astype(np.int64) - 1, 0, N)
    hi = np.clip(np.ceil(c + r).astype(np.int64) + 2, 0, N)
    rmin = r.min(axis=1)
    SL = 32

    def slab(z0):
        z1 = min(z0 + SL, N)
        f = np.full((z1 - z0, N, N), np.inf, np.float32)       # [z, y, x]
        which = np.zeros((z1 - z0, N, N), np.uint8)
        for i in range(blobs):
            a, b = max(z0, int(lo[i, 2])), min(z1, int(hi[i, 2]))
            if a >= b:
                continue
            ys, xs = slice(int(lo[i, 1]), int(hi[i, 1])), slice(int(lo[i, 0]), int(hi[i, 0]))
            zz = np.arange(a, b, dtype=np.float32)[:, None, None]
            yy = np.arange(ys.start, ys.stop, dtype=np.float32)[None, :, None]
            xx = np.arange(xs.start, xs.stop, dtype=np.float32)[None, None, :]
            d = ((xx - c[i, 0]) / r[i, 0]) ** 2 + ((yy - c[i, 1]) / r[i, 1]) ** 2 + ((zz - c[i, 2]) / r[i, 2]) ** 2
            fs, ws = f[a - z0:b - z0, ys, xs], which[a - z0:b - z0, ys, xs]
            upd = d < fs
            fs[upd] = d[upd]
            ws[upd] = i
        inside = f < 1.0
        shell = inside & (f > ((1.0 - 2.5 / rmin) ** 2).astype(np.float32)[which])
        zi, yi, xi = np.nonzero(inside)
        zi = zi + z0
        sh = shell[inside]
        struts = (((xi & 7) == 0) & ((yi & 7) == 0)) | (((yi & 7) == 0) & ((zi & 7) == 0)) | (((xi & 7) == 0) & ((zi & 7) == 0))
        keep = sh | struts
        w = which[inside][keep]
        return np.stack([xi[keep], yi[keep], zi[keep]], 1).astype(np.int32), (w.astype(np.uint32) % 16) + 1

    starts = list(range(0, N, SL))
    if len(starts) > 1:
        from concurrent.futures import ThreadPoolExecutor
        import os
        with ThreadPoolExecutor(min(len(starts), os.cpu_count() or 1, 16)) as ex:
            parts = list(ex.map(slab, starts))
    else:
        parts = [slab(0)]
    return np.concatenate([p[0] for p in parts]), np.concatenate([p[1] for p in parts]), default_materials(16)
