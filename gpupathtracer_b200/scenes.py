"""Procedural stand-ins for the scenes BASELINE.json names (no scene asset ships with the reference:
scene.txt:1 points at res/objects/22323.obj and there is no res/ directory).

Every generator is deterministic in its seed and returns a `SceneData` whose arrays are exactly what
the OBJ path of the reference would hand to Scene::LoadScene (Scene.cpp:233-292): per-triangle vertices
(position, vertex normal, uv), per-triangle material id (matId = 2*k, Scene.cpp:153-158) and the
material list with the sky at k = 0 (Scene.cpp:236-241).
"""
from dataclasses import dataclass, field
import numpy as np


@dataclass
class SceneData:
    name: str
    verts: np.ndarray            # (n, 3, 8) float32: pos.xyz, normal.xyz, uv.xy
    mat_ids: np.ndarray          # (n,) uint32, = 2*k
    materials: np.ndarray        # (m, 12) float32 view of pt_material records (ints bit-cast)
    env_color: tuple = (1.0, 1.0, 1.0)
    camera: dict = field(default_factory=dict)   # pos, rot(yaw,pitch,roll), fov, focus, aperture
    textures: list = field(default_factory=list)  # list of (h, w, 4) float32 arrays

    @property
    def n_tris(self):
        return int(self.verts.shape[0])


def _i2f(i):
    return np.array([i], dtype=np.int32).view(np.float32)[0]


def material(albedo, roughness=1.0, metallic=0.0, emission=(0, 0, 0), albedo_tex=-1, props_tex=-1):
    """One pt_material record (include/b200pt.h). `roughness` is the stored props.g; the shader squares it
    (Material.glsl:47). isEmissive as Scene.cpp:109."""
    e = np.asarray(emission, dtype=np.float32)
    rec = np.zeros(12, dtype=np.float32)
    rec[0:3] = albedo
    rec[3] = _i2f(albedo_tex)
    rec[4:7] = e
    rec[7] = _i2f(1 if float(e.sum()) > 1e-5 else 0)
    rec[8:11] = (0.0, roughness, metallic)
    rec[11] = _i2f(props_tex)
    return rec


def sky_material():
    # Scene.cpp:236-241: emission 25*(30,26,19), isEmissive = true; never sampled as a light (no triangles use it)
    rec = material((0, 0, 0), emission=(750.0, 650.0, 475.0))
    return rec


class _Builder:
    def __init__(self):
        self.v = []
        self.m = []

    def add(self, pos, nrm, uv, mat_k):
        """pos/nrm: (n,3,3), uv: (n,3,2)."""
        n = pos.shape[0]
        rec = np.zeros((n, 3, 8), dtype=np.float32)
        rec[:, :, 0:3] = pos
        rec[:, :, 3:6] = nrm
        rec[:, :, 6:8] = uv
        self.v.append(rec)
        self.m.append(np.full(n, 2 * mat_k, dtype=np.uint32))

    def quad(self, p00, p10, p11, p01, normal, mat_k, nu=1, nv=1):
        p00, p10, p11, p01 = (np.asarray(p, dtype=np.float64) for p in (p00, p10, p11, p01))
        u = np.linspace(0, 1, nu + 1)
        v = np.linspace(0, 1, nv + 1)
        U, V = np.meshgrid(u, v, indexing="ij")
        P = ((1 - U)[..., None] * (1 - V)[..., None] * p00 + U[..., None] * (1 - V)[..., None] * p10 +
             U[..., None] * V[..., None] * p11 + (1 - U)[..., None] * V[..., None] * p01)
        N = np.broadcast_to(np.asarray(normal, dtype=np.float64), P.shape)
        T = np.stack([U, V], axis=-1)
        self.grid(P, N, T, mat_k)

    def grid(self, P, N, T, mat_k, flip=False):
        """P,N: (a,b,3), T: (a,b,2) -> 2*(a-1)*(b-1) triangles."""
        a, b = P.shape[0], P.shape[1]
        i, j = np.meshgrid(np.arange(a - 1), np.arange(b - 1), indexing="ij")
        i = i.ravel(); j = j.ravel()
        idx = [(i, j), (i + 1, j), (i + 1, j + 1), (i, j + 1)]
        tri_a = [idx[0], idx[1], idx[2]]
        tri_b = [idx[0], idx[2], idx[3]]
        if flip:
            tri_a = tri_a[::-1]; tri_b = tri_b[::-1]
        for tri in (tri_a, tri_b):
            pos = np.stack([P[c[0], c[1]] for c in tri], axis=1)
            nrm = np.stack([N[c[0], c[1]] for c in tri], axis=1)
            uv = np.stack([T[c[0], c[1]] for c in tri], axis=1)
            self.add(pos.astype(np.float32), nrm.astype(np.float32), uv.astype(np.float32), mat_k)

    def sphere(self, center, radius, nu, nv, mat_k, disp=None, inward=False):
        th = np.linspace(0.0, np.pi, nv + 1)
        ph = np.linspace(0.0, 2 * np.pi, nu + 1)
        TH, PH = np.meshgrid(th, ph, indexing="ij")
        D = np.stack([np.sin(TH) * np.cos(PH), np.cos(TH), np.sin(TH) * np.sin(PH)], axis=-1)
        r = radius * (1.0 + (disp(D) if disp is not None else 0.0))
        P = np.asarray(center) + D * np.asarray(r)[..., None]
        N = -D if inward else D
        T = np.stack([PH / (2 * np.pi), TH / np.pi], axis=-1)
        self.grid(P, N, T, mat_k)

    def cylinder(self, base, radius, height, nu, nv, mat_k, axis=1):
        ph = np.linspace(0.0, 2 * np.pi, nu + 1)
        h = np.linspace(0.0, height, nv + 1)
        H, PH = np.meshgrid(h, ph, indexing="ij")
        c, s = np.cos(PH), np.sin(PH)
        comps = [radius * c, H, radius * s]
        ncomps = [c, np.zeros_like(c), s]
        if axis == 0:
            comps = [H, radius * c, radius * s]; ncomps = [np.zeros_like(c), c, s]
        elif axis == 2:
            comps = [radius * c, radius * s, H]; ncomps = [c, s, np.zeros_like(c)]
        P = np.asarray(base) + np.stack(comps, axis=-1)
        N = np.stack(ncomps, axis=-1)
        T = np.stack([PH / (2 * np.pi), H / max(height, 1e-6)], axis=-1)
        self.grid(P, N, T, mat_k)

    def box_inward(self, lo, hi, mats, n=1):
        """Room: 6 faces with inward normals. mats: dict face->k or single k."""
        lo = np.asarray(lo, dtype=np.float64); hi = np.asarray(hi, dtype=np.float64)
        x0, y0, z0 = lo; x1, y1, z1 = hi
        get = (lambda f: mats[f]) if isinstance(mats, dict) else (lambda f: mats)
        self.quad((x0, y0, z0), (x1, y0, z0), (x1, y0, z1), (x0, y0, z1), (0, 1, 0), get("floor"), n, n)
        self.quad((x0, y1, z0), (x0, y1, z1), (x1, y1, z1), (x1, y1, z0), (0, -1, 0), get("ceil"), n, n)
        self.quad((x0, y0, z0), (x0, y0, z1), (x0, y1, z1), (x0, y1, z0), (1, 0, 0), get("left"), n, n)
        self.quad((x1, y0, z0), (x1, y1, z0), (x1, y1, z1), (x1, y0, z1), (-1, 0, 0), get("right"), n, n)
        self.quad((x0, y0, z0), (x0, y1, z0), (x1, y1, z0), (x1, y0, z0), (0, 0, 1), get("back"), n, n)
        self.quad((x0, y0, z1), (x1, y0, z1), (x1, y1, z1), (x0, y1, z1), (0, 0, -1), get("front"), n, n)

    def finish(self, name, materials, env_color, camera, textures=None):
        verts = np.ascontiguousarray(np.concatenate(self.v, axis=0))
        mats = np.ascontiguousarray(np.concatenate(self.m, axis=0))
        return SceneData(name, verts, mats, np.ascontiguousarray(np.stack(materials).astype(np.float32)),
                         tuple(env_color), camera, textures or [])


def _noise_disp(seed, amp, freq=3.0, octaves=3):
    rng = np.random.default_rng(seed)
    ks = [rng.normal(size=(3,)) * freq * (2 ** o) for o in range(octaves)]
    phs = [rng.uniform(0, 2 * np.pi) for _ in range(octaves)]

    def f(D):
        out = np.zeros(D.shape[:-1])
        for o, (k, p) in enumerate(zip(ks, phs)):
            out += (amp / (2 ** o)) * np.sin(D @ k + p)
        return out
    return f


def cornell(n=1, seed=0, light_emission=(17.0, 12.0, 4.0)):
    """Small closed room with two blocks-as-spheres and a ceiling light: the unit-test scene."""
    b = _Builder()
    mats = [sky_material(),
            material((0.73, 0.73, 0.73), 1.0, 0.0),            # 1 white
            material((0.65, 0.05, 0.05), 1.0, 0.0),            # 2 red
            material((0.12, 0.45, 0.15), 1.0, 0.0),            # 3 green
            material((0.0, 0.0, 0.0), 1.0, 0.0, light_emission),   # 4 light
            material((0.9, 0.9, 0.9), 0.35, 1.0),              # 5 glossy metal
            material((0.2, 0.3, 0.8), 0.6, 0.0)]               # 6 blue plastic
    b.box_inward((-1, 0, -1), (1, 2, 1), {"floor": 1, "ceil": 1, "back": 1, "front": 1, "left": 2, "right": 3}, n)
    b.quad((-0.3, 1.995, -0.3), (-0.3, 1.995, 0.3), (0.3, 1.995, 0.3), (0.3, 1.995, -0.3), (0, -1, 0), 4, max(1, n // 2), max(1, n // 2))
    b.sphere((-0.4, 0.4, -0.3), 0.4, 8 * n, 6 * n, 5)
    b.sphere((0.45, 0.3, 0.3), 0.3, 8 * n, 6 * n, 6, disp=_noise_disp(seed + 1, 0.08))
    cam = dict(pos=(0.0, 1.0, 0.95), rot=(0.0, 0.0, 0.0), fov=np.radians(60.0), focus=1.5, aperture=0.0)
    return b.finish("cornell", mats, (0.0, 0.0, 0.0), cam)


def bunny_like(seed=1):
    """Config 1: 'Stanford bunny (~69k tris)': noise-displaced UV sphere 186 x 186 x 2 = 69 192 triangles on a
    ground quad with two emissive quads."""
    b = _Builder()
    mats = [sky_material(),
            material((0.8, 0.75, 0.7), 0.9, 0.0),
            material((0.5, 0.5, 0.5), 1.0, 0.0),
            material((0, 0, 0), 1.0, 0.0, (20.0, 18.0, 15.0))]
    b.sphere((0, 1.0, 0), 1.0, 186, 186, 1, disp=_noise_disp(seed, 0.12, 2.5, 4))
    b.quad((-6, -0.2, -6), (-6, -0.2, 6), (6, -0.2, 6), (6, -0.2, -6), (0, 1, 0), 2)
    b.quad((-2, 4, -1), (-2, 4, 1), (-1, 4.5, 1), (-1, 4.5, -1), (0.4, -0.9, 0), 3)
    b.quad((1.5, 3.5, -1), (1.5, 3.5, 1), (2.5, 3.0, 1), (2.5, 3.0, -1), (-0.4, -0.9, 0), 3)
    cam = dict(pos=(0.0, 1.2, 4.5), rot=(0.0, 0.0, 0.0), fov=np.radians(45.0), focus=4.5, aperture=0.0)
    return b.finish("bunny_like", mats, (1.0, 1.0, 1.0), cam)


def atrium(target_tris=262144, seed=2, aperture=0.0, glossy=False):
    """Config 2 (and 3 at a higher count): 'Sponza-like' atrium: a hall with two colonnade tiers, arches, displaced
    drapes and emissive ceiling strips. target_tris = 262144 gives exactly 262 144 triangles; other targets scale
    the tessellation and land within a few percent."""
    rng = np.random.default_rng(seed)
    b = _Builder()
    r0 = 0.25 if glossy else 0.8
    mats = [sky_material()]
    diffuse = [(0.75, 0.7, 0.6), (0.6, 0.55, 0.5), (0.7, 0.3, 0.25), (0.3, 0.45, 0.65), (0.35, 0.6, 0.3),
               (0.8, 0.8, 0.8), (0.5, 0.4, 0.3), (0.65, 0.6, 0.2)]
    for c in diffuse:
        mats.append(material(c, max(r0, 0.5) if not glossy else 0.45, 0.0))        # 1..8
    for c, r in [((0.95, 0.93, 0.88), 0.3), ((0.9, 0.6, 0.3), 0.4), ((0.7, 0.7, 0.75), 0.25), ((0.4, 0.4, 0.45), 0.5)]:
        mats.append(material(c, r if not glossy else r * 0.6, 1.0))                # 9..12
    mats.append(material((0, 0, 0), 1.0, 0.0, (30.0, 27.0, 22.0)))                # 13 emitter
    EMIT = 13
    L, W, H = 30.0, 12.0, 14.0
    # budget: hall shell ~2%, columns ~45%, arches ~15%, drapes ~36%, lights fixed
    s = np.sqrt(target_tris / 262144.0)
    nshell = max(2, int(round(24 * s)))
    b.box_inward((-L / 2, 0, -W / 2), (L / 2, H, W / 2),
                 {"floor": 1, "ceil": 6, "back": 2, "front": 2, "left": 7, "right": 7}, nshell)
    # emissive ceiling strips: 32 quads = 64 triangles
    for k in range(32):
        x0 = -L / 2 + 1.0 + (L - 2.0) * k / 32.0
        b.quad((x0, H - 0.02, -0.6), (x0, H - 0.02, 0.6), (x0 + 0.5, H - 0.02, 0.6), (x0 + 0.5, H - 0.02, -0.6), (0, -1, 0), EMIT)
    # two tiers of columns along both long sides
    ncol = 14
    cu = max(6, int(round(34 * s))); cv = max(4, int(round(30 * s)))
    for tier, (y0, hgt, rad) in enumerate([(0.0, 6.0, 0.35), (6.6, 5.0, 0.28)]):
        for side in (-1, 1):
            for k in range(ncol):
                x = -L / 2 + 2.0 + (L - 4.0) * k / (ncol - 1)
                z = side * (W / 2 - 2.0)
                b.cylinder((x, y0, z), rad, hgt, cu, cv, 9 + (k + tier) % 4 if k % 3 == 0 else 1 + (k + tier) % 8)
    # gallery floor slabs between tiers
    for side in (-1, 1):
        z0, z1 = (side * (W / 2 - 2.6), side * W / 2) if side > 0 else (side * W / 2, side * (W / 2 - 2.6))
        b.quad((-L / 2, 6.3, z0), (-L / 2, 6.3, z1), (L / 2, 6.3, z1), (L / 2, 6.3, z0), (0, 1, 0), 2, nshell, 2)
        b.quad((-L / 2, 6.2, z0), (L / 2, 6.2, z0), (L / 2, 6.2, z1), (-L / 2, 6.2, z1), (0, -1, 0), 2, nshell, 2)
    # arches: half-tori segments between neighbouring columns (as swept half circles)
    au = max(4, int(round(36 * s))); av = max(3, int(round(16 * s)))
    for side in (-1, 1):
        z = side * (W / 2 - 2.0)
        for k in range(ncol - 1):
            xa = -L / 2 + 2.0 + (L - 4.0) * k / (ncol - 1)
            xb = -L / 2 + 2.0 + (L - 4.0) * (k + 1) / (ncol - 1)
            cx, R = 0.5 * (xa + xb), 0.5 * (xb - xa)
            t = np.linspace(0, np.pi, au + 1)
            a = np.linspace(0, 2 * np.pi, av + 1)
            Tt, Aa = np.meshgrid(t, a, indexing="ij")
            r = 0.18
            P = np.stack([cx + (R + r * np.cos(Aa)) * np.cos(Tt), 6.0 - 0.6 + (R + r * np.cos(Aa)) * np.sin(Tt) * 0.45,
                          z + r * np.sin(Aa)], axis=-1)
            N = np.stack([np.cos(Aa) * np.cos(Tt), np.cos(Aa) * np.sin(Tt), np.sin(Aa)], axis=-1)
            b.grid(P, N, np.stack([Tt / np.pi, Aa / (2 * np.pi)], axis=-1), 5)
    # drapes: displaced vertical grids hanging across the hall
    ndrape = 6
    du = max(8, int(round(96 * s))); dv = max(8, int(round(96 * s)))
    for k in range(ndrape):
        x = -L / 2 + 3.0 + (L - 6.0) * k / (ndrape - 1)
        zz = np.linspace(-W / 2 + 2.8, W / 2 - 2.8, du + 1)
        yy = np.linspace(7.0, 12.5, dv + 1)
        Z, Y = np.meshgrid(zz, yy, indexing="ij")
        ph = rng.uniform(0, 2 * np.pi, size=3)
        X = x + 0.25 * np.sin(1.7 * Z + ph[0]) + 0.12 * np.sin(4.1 * Z + 2.3 * Y + ph[1]) + 0.05 * np.sin(9.0 * Y + ph[2])
        P = np.stack([X, Y, Z], axis=-1)
        N = np.broadcast_to(np.array([1.0, 0.0, 0.0]), P.shape)
        b.grid(P, N, np.stack([(Z + W / 2) / W, (Y - 7.0) / 5.5], axis=-1), 3 + k % 3)
    cam = dict(pos=(-L / 2 + 1.5, 3.0, 0.5), rot=(np.pi / 2 + 0.05, 0.12, 0.0), fov=np.radians(55.0), focus=9.0,
               aperture=aperture)
    return b.finish("atrium", mats, (1.0, 1.0, 1.0), cam)


def shells(n_shells=8, tris_per_shell=1250000, seed=4):
    """Config 4: nested noise-displaced shells (depth complexity >= n_shells); 8 x 1.25 M = 10 M triangles."""
    b = _Builder()
    mats = [sky_material(), material((0.7, 0.7, 0.7), 0.8, 0.0), material((0, 0, 0), 1.0, 0.0, (10, 10, 10))]
    n = int(round(np.sqrt(tris_per_shell / 2.0)))
    for k in range(n_shells):
        b.sphere((0, 0, 0), 1.0 + 0.35 * k, n, n, 1 if k else 2, disp=_noise_disp(seed + k, 0.05, 3.0 + k, 3))
    cam = dict(pos=(0.0, 0.0, 6.0), rot=(0.0, 0.0, 0.0), fov=np.radians(45.0), focus=6.0, aperture=0.0)
    return b.finish("shells", mats, (1.0, 1.0, 1.0), cam)


def random_soup(n=2000, seed=7, extent=4.0, size=0.35):
    """Unstructured triangle soup with a few emitters: stresses overlapping boxes / spatial splits."""
    rng = np.random.default_rng(seed)
    c = rng.uniform(-extent, extent, size=(n, 1, 3))
    pos = (c + rng.normal(scale=size, size=(n, 3, 3)) * rng.uniform(0.2, 2.0, size=(n, 1, 1))).astype(np.float32)
    e1 = pos[:, 1] - pos[:, 0]; e2 = pos[:, 2] - pos[:, 0]
    nn = np.cross(e1, e2); nn /= np.maximum(np.linalg.norm(nn, axis=1, keepdims=True), 1e-20)
    nrm = np.repeat(nn[:, None, :], 3, axis=1)
    uv = rng.uniform(0, 1, size=(n, 3, 2)).astype(np.float32)
    b = _Builder()
    mats = [sky_material(), material((0.7, 0.6, 0.5), 0.7, 0.0), material((0.9, 0.9, 0.9), 0.3, 1.0),
            material((0, 0, 0), 1.0, 0.0, (8.0, 8.0, 8.0))]
    k = rng.integers(1, 3, size=n)
    k[rng.choice(n, size=max(2, n // 50), replace=False)] = 3
    for kk in (1, 2, 3):
        sel = k == kk
        if sel.any():
            b.add(pos[sel], nrm[sel].astype(np.float32), uv[sel], kk)
    cam = dict(pos=(0.0, 0.0, 2.5 * extent), rot=(0.0, 0.0, 0.0), fov=np.radians(50.0), focus=2.5 * extent, aperture=0.0)
    return b.finish("random_soup", mats, (0.5, 0.6, 0.8), cam)


def random_rays(n, lo, hi, seed=0xB200, tmax=1e20, any_hit=False):
    """Config 4 ray batch: origins uniform in the AABB, directions uniform on the sphere."""
    rng = np.random.default_rng(seed)
    lo = np.asarray(lo, dtype=np.float32); hi = np.asarray(hi, dtype=np.float32)
    rays = np.zeros((n, 8), dtype=np.float32)
    rays[:, 0:3] = rng.uniform(0, 1, size=(n, 3)).astype(np.float32) * (hi - lo) + lo
    z = rng.uniform(-1, 1, size=n); ph = rng.uniform(0, 2 * np.pi, size=n)
    r = np.sqrt(np.maximum(0.0, 1 - z * z))
    rays[:, 4] = r * np.cos(ph); rays[:, 5] = r * np.sin(ph); rays[:, 6] = z
    if any_hit:
        diag = float(np.linalg.norm(hi - lo))
        rays[:, 7] = rng.uniform(0, diag, size=n).astype(np.float32)
    else:
        rays[:, 7] = tmax
    return rays
