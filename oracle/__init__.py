"""oracle -- TEST INFRASTRUCTURE ONLY: ctypes bindings of the CPU checkers.

  oracle/libpt_oracle.so      our CPU restatement of the reference's GLSL path (oracle/pt_oracle.cpp)
  oracle/_ref/libref_shim.so  the reference's OWN C++ (SBVH builder, AABB, Camera, HybridTaus, tonemap)
                              compiled in place from /root/reference by oracle/Makefile

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(_HERE, "libpt_oracle.so")
REF_SO = os.path.join(_HERE, "_ref", "libref_shim.so")
REFERENCE_ROOT = "/root/reference"


def build(ref=True):
    subprocess.run(["make", "-C", _HERE, "oracle"], check=True, capture_output=True)
    if ref and os.path.isdir(REFERENCE_ROOT):
        subprocess.run(["make", "-C", _HERE, "ref"], check=True, capture_output=True)


class OrTexture(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("rgba", C.POINTER(C.c_float))]


class OrScene(C.Structure):
    _fields_ = [("nodes", C.c_void_p), ("n_nodes", C.c_int), ("refs", C.c_void_p), ("n_refs", C.c_int),
                ("tris", C.c_void_p), ("n_tris", C.c_int), ("materials", C.c_void_p), ("n_materials", C.c_int),
                ("textures", C.POINTER(OrTexture)), ("n_textures", C.c_int), ("lights", C.c_void_p),
                ("n_lights", C.c_int), ("total_light_area", C.c_float), ("env_color", C.c_float * 3),
                ("env_faces", C.c_void_p), ("env_size", C.c_int)]


class OrCamera(C.Structure):
    _fields_ = [("lower_left", C.c_float * 3), ("horizontal", C.c_float * 3), ("vertical", C.c_float * 3),
                ("position", C.c_float * 3), ("u", C.c_float * 3), ("v", C.c_float * 3), ("lens_radius", C.c_float)]


class OrConfig(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("max_depth", C.c_int), ("microfacet_model", C.c_int),
                ("n_threads", C.c_int), ("sun_light", C.c_int), ("sun_dir", C.c_float * 3), ("sun_radius", C.c_float)]


class OrCounters(C.Structure):
    _fields_ = [("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("node_pairs", C.c_uint64),
                ("leaf_visits", C.c_uint64), ("tri_tests", C.c_uint64), ("paths", C.c_uint64),
                ("shadow_node_pairs", C.c_uint64), ("shadow_tri_tests", C.c_uint64), ("null_shadow_rays", C.c_uint64)]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_SO):
            build(ref=False)
        _lib = C.CDLL(ORACLE_SO)
        _lib.or_tonemap.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        _lib.or_slab_test.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p]
        _lib.or_slab_test.restype = C.c_int
        _lib.or_image_plane.argtypes = [C.c_float] * 4 + [C.c_void_p, C.c_void_p, C.POINTER(OrCamera)]
        _lib.or_hybrid_taus.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        _lib.or_trace_closest.argtypes = [C.POINTER(OrScene), C.c_void_p, C.c_int64, C.c_void_p, C.POINTER(OrCounters), C.c_int]
        _lib.or_trace_any.argtypes = [C.POINTER(OrScene), C.c_void_p, C.c_int64, C.c_void_p, C.POINTER(OrCounters), C.c_int]
        _lib.or_trace_closest_bruteforce.argtypes = [C.POINTER(OrScene), C.c_void_p, C.c_int64, C.c_void_p, C.c_int]
        _lib.or_render.argtypes = [C.POINTER(OrScene), C.POINTER(OrCamera), C.POINTER(OrConfig), C.c_void_p, C.c_void_p,
                                   C.c_int, C.POINTER(OrCounters)]
        _lib.or_sample_environment.argtypes = [C.POINTER(OrScene), C.c_void_p, C.c_int64, C.c_void_p]
        _lib.or_equirect_to_cube.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        _lib.or_raygen.argtypes = [C.POINTER(OrCamera), C.POINTER(OrConfig), C.c_float, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]
    return _lib


def have_ref():
    return os.path.exists(REF_SO)


# ---- parity libm (gpupathtracer_b200/csrc/pt_parity_libm.h compiled into every checker)
PLIBM_FNS = ["sin", "cos", "log", "exp", "pow"]


def set_parity_libm(on):
    """GLSL's sin / cos / log / exp / pow inside libpt_oracle.so: glibc (False) or pt_parity_libm.h (True)."""
    lib().or_set_parity_libm(1 if on else 0)


def parity_libm_probe(fn, x, y=None):
    """pt_parity_libm.h as the host compiler evaluates it, element by element."""
    l = lib()
    l.or_parity_libm_probe.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
    x = np.ascontiguousarray(x, dtype=np.float32)
    yy = np.ascontiguousarray(y if y is not None else np.zeros_like(x), dtype=np.float32)
    out = np.zeros_like(x)
    l.or_parity_libm_probe(PLIBM_FNS.index(fn), _p(x), _p(yy), x.size, _p(out))
    return out


def ref():
    """The reference's own code; None when oracle/_ref has not been built (e.g. no /root/reference)."""
    global _ref
    if _ref is None:
        if not os.path.exists(REF_SO):
            return None
        _ref = C.CDLL(REF_SO)
        _ref.ref_build_sbvh.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.c_int),
                                        C.POINTER(C.POINTER(C.c_int)), C.POINTER(C.c_int)]
        _ref.ref_free.argtypes = [C.c_void_p]
        _ref.ref_camera_image_plane.argtypes = [C.c_float] * 4 + [C.c_void_p, C.c_void_p, C.c_void_p]
        _ref.ref_aabb_intersect.argtypes = [C.c_void_p] * 4 + [C.c_float, C.c_void_p]
        _ref.ref_aabb_intersect.restype = C.c_int
        _ref.ref_aabb_surface_area.argtypes = [C.c_void_p, C.c_void_p]
        _ref.ref_aabb_surface_area.restype = C.c_float
        _ref.ref_hybrid_taus.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        _ref.ref_tonemap_uncharted2.argtypes = [C.c_void_p, C.c_void_p]
    return _ref


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


GLSL_SO = os.path.join(_HERE, "_ref", "libref_glsl.so")
_glsl = None


def have_glsl():
    return os.path.exists(GLSL_SO)


def shade_probe_inputs(n, seed):
    """Random cases for or_shade_probe / ref_glsl_shade_probe: n x 16 floats {albedo.rgb, props.g (sqrt-roughness), props.b
    (metallic), normal, outgoing, light direction} + n x 4 RNG words. Views above and (10 %) below the surface, light directions
    over the whole sphere, rough to near-mirror, dielectric to metal."""
    g = np.random.default_rng(seed)
    def unit(m):
        v = g.normal(size=(m, 3)); return (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(np.float32)
    a = np.zeros((n, 16), np.float32)
    a[:, 0:3] = g.uniform(0.02, 1.0, size=(n, 3))
    a[:, 3] = g.uniform(0.02, 1.0, size=n) ** 1.5
    a[:, 4] = np.where(g.uniform(size=n) < 0.3, 1.0, np.where(g.uniform(size=n) < 0.5, 0.0, g.uniform(size=n)))
    nrm = unit(n); nrm[: n // 8] = (0.0, 1.0, 0.0)                       # the ConstructTBN branch for |n.y| > 0.99
    o = unit(n); flip = (np.sum(o * nrm, axis=1) < 0) & (g.uniform(size=n) < 0.9); o[flip] *= -1
    a[:, 5:8], a[:, 8:11], a[:, 11:14] = nrm, o, unit(n)
    rng = g.integers(129, 2 ** 32, size=(n, 4), dtype=np.uint64).astype(np.uint32)
    return a, rng


def shade_probe(inputs, rng, model=0):
    """or_shade_probe: the oracle's shading arithmetic on explicit inputs -> (n x 20 floats, RNG words after the sample)."""
    l = lib()
    l.or_shade_probe.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p]
    a = np.ascontiguousarray(inputs, dtype=np.float32); r = np.ascontiguousarray(rng, dtype=np.uint32).copy()
    out = np.zeros((a.shape[0], 20), np.float32)
    l.or_shade_probe(_p(a), _p(r), a.shape[0], model, _p(out))
    return out, r


def ref_glsl_shade_probe(inputs, rng, model=0):
    """The same probe through the reference's own shader functions compiled as C++ (oracle/ref_glsl.cpp); model 1 = the copy
    whose three model macros (Microfacet.glsl:88-90) point at the file's Beckmann functions; model 2 = that copy with scalar
    min / max meaning IEEE minNum / maxNum (what NVIDIA's FMNMX executes) instead of glm's ternaries. ref_glsl_light_probe and
    Oracle.ref_glsl_render take the same codes."""
    g = C.CDLL(GLSL_SO)
    a = np.ascontiguousarray(inputs, dtype=np.float32); r = np.ascontiguousarray(rng, dtype=np.uint32).copy()
    out = np.zeros((a.shape[0], 20), np.float32)
    if model in (4, 6):                                        # 0 | 4, 2 | 4: the builds with pt_parity_libm.h's transcendentals
        g.ref_glsl_shade_probe_model.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p]
        g.ref_glsl_shade_probe_model(_p(a), _p(r), a.shape[0], model, _p(out))
        return out, r
    fn = {0: g.ref_glsl_shade_probe, 1: g.ref_glsl_shade_probe_beckmann, 2: g.ref_glsl_shade_probe_beckmann_maxnum}[model]
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
    fn(_p(a), _p(r), a.shape[0], _p(out))
    return out, r


def light_probe_inputs(scene, n, seed):
    """Random cases for the light-sampling probes: surface points inside the scene's bounds with random normals, views mostly
    above the surface, the materials of shade_probe_inputs, random throughput; n x 20 floats + n x 4 RNG words."""
    a16, rng = shade_probe_inputs(n, seed)
    g = np.random.default_rng(seed + 1000)
    lo, hi = scene.bounds()
    a = np.zeros((n, 20), np.float32)
    a[:, 0:3] = g.uniform(lo, hi, size=(n, 3))
    a[:, 3:6], a[:, 6:9] = a16[:, 5:8], a16[:, 8:11]
    a[:, 9:12], a[:, 12], a[:, 13] = a16[:, 0:3], a16[:, 3], a16[:, 4]
    a[:, 14:17] = g.uniform(0.0, 2.0, size=(n, 3))
    return a, rng


def ref_glsl_light_probe(scene, inputs, rng, model=0):
    """GenerateLightSample + RandomLightVertex of Iterative.comp compiled as C++ (oracle/ref_glsl.cpp), on the scene's light
    table and triangles."""
    g = C.CDLL(GLSL_SO)
    g.ref_glsl_bind.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
    g.ref_glsl_bind_lights.argtypes = [C.c_void_p, C.c_int, C.c_float]
    g.ref_glsl_light_probe.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p]
    nodes = np.ascontiguousarray(scene.nodes, np.float32); refs = np.ascontiguousarray(scene.refs, np.int32)
    tris = np.ascontiguousarray(scene.tris, np.float32); lights = np.ascontiguousarray(scene.lights, np.float32)
    g.ref_glsl_bind(_p(nodes), _p(refs), _p(tris), tris.shape[0])
    g.ref_glsl_bind_lights(_p(lights), lights.shape[0], float(scene.total_light_area))
    a = np.ascontiguousarray(inputs, dtype=np.float32); r = np.ascontiguousarray(rng, dtype=np.uint32).copy()
    out = np.zeros((a.shape[0], 12), np.float32)
    g.ref_glsl_light_probe(_p(a), _p(r), a.shape[0], model, _p(out))
    return out, r


def ref_glsl_raygen(cam_array, width, height, total_light_area, rng_table, first, n):
    """InitRay + RayIndexToMorton of Iterative.comp compiled as C++: n x 12 floats {pixel (int bits), origin, direction,
    throughput.x, bxdfPdf0, bxdfPdf1, neePdf}; the pixels' RNG words in rng_table are advanced in place."""
    g = C.CDLL(GLSL_SO)
    g.ref_glsl_raygen_probe.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]
    cam = np.ascontiguousarray(cam_array, dtype=np.float32)
    out = np.zeros((n, 12), np.float32)
    g.ref_glsl_raygen_probe(_p(cam), width, height, float(total_light_area), _p(rng_table), first, n, _p(out))
    return out


def ref_glsl_present(accum, num_samples, exposure=1.68, parity_libm=False):
    """Present.frag compiled as C++ (oracle/ref_glsl.cpp) on an (h, w, 4) accumulation image -> (h, w, 3) float32, the colour the
    fragment shader writes; exposure is the uniform the reference sets to kExposure (Renderer.cpp:18). parity_libm: the build
    whose pow() is pt_parity_libm.h's."""
    g = C.CDLL(GLSL_SO)
    fn = g.ref_glsl_present_pl if parity_libm else g.ref_glsl_present
    fn.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_void_p]
    a = np.ascontiguousarray(accum, dtype=np.float32)
    out = np.zeros(a.shape[:2] + (3,), np.float32)
    fn(_p(a), a.shape[1], a.shape[0], int(num_samples), float(exposure), _p(out))
    return out


def equirect_uv(sample_coords):
    """or_equirect_uv: the oracle's statement of EquirectangularConverter.frag's coordinate arithmetic, (n, 3) -> (n, 2)."""
    p = np.ascontiguousarray(sample_coords, dtype=np.float32); out = np.zeros((p.shape[0], 2), np.float32)
    lib().or_equirect_uv.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
    lib().or_equirect_uv(_p(p), p.shape[0], _p(out))
    return out


def ref_glsl_equirect_uv(sample_coords):
    """SampleSphericalMap(normalize(sampleCoords)) of EquirectangularConverter.frag compiled as C++."""
    g = C.CDLL(GLSL_SO)
    g.ref_glsl_equirect_uv.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
    p = np.ascontiguousarray(sample_coords, dtype=np.float32); out = np.zeros((p.shape[0], 2), np.float32)
    g.ref_glsl_equirect_uv(_p(p), p.shape[0], _p(out))
    return out


def ref_glsl_hybrid_taus(state4, n):
    g = C.CDLL(GLSL_SO)
    g.ref_glsl_hybrid_taus.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    st = np.ascontiguousarray(state4, dtype=np.uint32).copy(); out = np.zeros(n, np.float32)
    g.ref_glsl_hybrid_taus(_p(st), n, _p(out))
    return out, st


class RefGlsl:
    """The reference's own traversal SHADER SOURCE (IfIfClosestHit / IfIfAnyHit and what they call, common/BVH.glsl +
    common/Geometry.glsl) compiled as C++ in place by oracle/Makefile (oracle/_ref/libref_glsl.so, harness oracle/ref_glsl.cpp).
    One binding at a time: the samplers are globals, as uniforms are."""

    def __init__(self, nodes, refs, tris_pre):
        global _glsl
        if _glsl is None:
            _glsl = C.CDLL(GLSL_SO)
            _glsl.ref_glsl_bind.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
            _glsl.ref_glsl_trace_closest.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int]
            _glsl.ref_glsl_trace_any.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int]
        self.nodes = np.ascontiguousarray(nodes, dtype=np.float32)
        self.refs = np.ascontiguousarray(refs, dtype=np.int32)
        self.tris = np.ascontiguousarray(tris_pre, dtype=np.float32)

    def _bind(self):
        _glsl.ref_glsl_bind(_p(self.nodes), _p(self.refs), _p(self.tris), self.tris.shape[0])

    def trace_closest(self, rays, n_threads=0):
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        hits = np.zeros((rays.shape[0], 4), dtype=np.float32)
        self._bind()
        _glsl.ref_glsl_trace_closest(_p(rays), rays.shape[0], _p(hits), n_threads)
        return hits

    def trace_any(self, rays, n_threads=0):
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        flags = np.zeros(rays.shape[0], dtype=np.uint8)
        self._bind()
        _glsl.ref_glsl_trace_any(_p(rays), rays.shape[0], _p(flags), n_threads)
        return flags


def ref_build_sbvh(tris_abs):
    """BoundingVolumeHierarchy::BuildBinnedSpatial (BVH.cpp:2197-2293) on absolute-position triangles.
    Returns (nodes (n,8) float32 in NewLayout, refs int32). stdout of the reference is silenced."""
    r = ref()
    assert r is not None, "oracle/_ref/libref_shim.so not built"
    t = np.ascontiguousarray(tris_abs, dtype=np.float32)
    nodes_p, refs_p = C.POINTER(C.c_float)(), C.POINTER(C.c_int)()
    nn, nr = C.c_int(), C.c_int()
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
    os.dup2(devnull, 1)
    try:
        rc = r.ref_build_sbvh(_p(t), t.shape[0], C.byref(nodes_p), C.byref(nn), C.byref(refs_p), C.byref(nr))
    finally:
        os.dup2(saved, 1)
        os.close(saved)
        os.close(devnull)
    assert rc == 0
    nodes = np.ctypeslib.as_array(nodes_p, shape=(nn.value, 8)).copy()
    refs = np.ctypeslib.as_array(refs_p, shape=(nr.value,)).astype(np.int32).copy()
    r.ref_free(nodes_p)
    r.ref_free(refs_p)
    return nodes, refs


class Oracle:
    """Scene tables + entry points of libpt_oracle.so. `scene` is a gpupathtracer_b200.Scene-like object
    exposing tris (pre-subtracted), materials, lights, total_light_area, env_color, textures, nodes, refs."""

    def __init__(self, scene, nodes=None, refs=None):
        self.l = lib()
        self.tris = np.ascontiguousarray(scene.tris, dtype=np.float32)
        self.materials = np.ascontiguousarray(scene.materials, dtype=np.float32)
        self.lights = np.ascontiguousarray(scene.lights, dtype=np.float32)
        self.nodes = np.ascontiguousarray(nodes if nodes is not None else scene.nodes, dtype=np.float32)
        self.refs = np.ascontiguousarray(refs if refs is not None else scene.refs, dtype=np.int32)
        self.textures = [np.ascontiguousarray(t, dtype=np.float32) for t in getattr(scene, "textures", [])]
        self.tex_arr = (OrTexture * max(1, len(self.textures)))()
        for i, t in enumerate(self.textures):
            self.tex_arr[i].width, self.tex_arr[i].height = t.shape[1], t.shape[0]
            self.tex_arr[i].rgba = t.ctypes.data_as(C.POINTER(C.c_float))
        s = OrScene()
        s.nodes, s.n_nodes = _p(self.nodes), self.nodes.shape[0]
        s.refs, s.n_refs = _p(self.refs), self.refs.shape[0]
        s.tris, s.n_tris = _p(self.tris), self.tris.shape[0]
        s.materials, s.n_materials = _p(self.materials), self.materials.shape[0]
        s.textures, s.n_textures = self.tex_arr, len(self.textures)
        s.lights, s.n_lights = _p(self.lights), self.lights.shape[0]
        s.total_light_area = float(scene.total_light_area)
        s.env_color = (C.c_float * 3)(*scene.env_color)
        cube = getattr(scene, "env_cubemap", None)
        if cube is not None:
            self.env_faces = np.ascontiguousarray(cube, dtype=np.float32)
            assert self.env_faces.shape[0] == 6 and self.env_faces.shape[1] == self.env_faces.shape[2]
            s.env_faces, s.env_size = _p(self.env_faces), self.env_faces.shape[1]
        else:
            s.env_faces, s.env_size = None, 0
        self.s = s

    def sample_environment(self, dirs):
        dirs = np.ascontiguousarray(dirs, dtype=np.float32)
        out = np.zeros((dirs.shape[0], 3), dtype=np.float32)
        self.l.or_sample_environment(C.byref(self.s), _p(dirs), C.c_int64(dirs.shape[0]), _p(out))
        return out

    def trace_closest(self, rays, n_threads=0):
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        hits = np.zeros((rays.shape[0], 4), dtype=np.float32)
        c = OrCounters()
        self.l.or_trace_closest(C.byref(self.s), _p(rays), rays.shape[0], _p(hits), C.byref(c), n_threads)
        return hits, c.as_dict()

    def trace_any(self, rays, n_threads=0):
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        flags = np.zeros(rays.shape[0], dtype=np.uint8)
        c = OrCounters()
        self.l.or_trace_any(C.byref(self.s), _p(rays), rays.shape[0], _p(flags), C.byref(c), n_threads)
        return flags, c.as_dict()

    def trace_ops(self, rays, any_hit=False):
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        n = rays.shape[0]
        self.l.or_trace_ops.restype = C.c_int64
        self.l.or_trace_ops.argtypes = [C.POINTER(OrScene), C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p]
        total = self.l.or_trace_ops(C.byref(self.s), _p(rays), n, int(any_hit), None, None)
        ops = np.zeros(total, dtype=np.uint8); off = np.zeros(n + 1, dtype=np.int64)
        self.l.or_trace_ops(C.byref(self.s), _p(rays), n, int(any_hit), _p(ops), _p(off))
        return ops, off

    def trace_closest_bruteforce(self, rays, n_threads=0):
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        hits = np.zeros((rays.shape[0], 4), dtype=np.float32)
        self.l.or_trace_closest_bruteforce(C.byref(self.s), _p(rays), rays.shape[0], _p(hits), n_threads)
        return hits

    @staticmethod
    def camera(cam_array):
        c = OrCamera()
        a = np.asarray(cam_array, dtype=np.float32)
        for i, name in enumerate(["lower_left", "horizontal", "vertical", "position", "u", "v"]):
            setattr(c, name, (C.c_float * 3)(*a[3 * i:3 * i + 3]))
        c.lens_radius = float(a[18])
        return c

    def render(self, cam_array, width, height, rng, accum, n_frames=1, max_depth=66, model=0, n_threads=0, sun=None,
               parity_libm=False):
        """In place on rng (px,4) uint32 and accum (h,w,4) float32. Returns counters. sun = (direction, radius) turns the
        shader's SUN_LIGHT branch on. parity_libm: sin / cos / log / exp from pt_parity_libm.h for this call."""
        set_parity_libm(parity_libm)
        try:
            return self._render(cam_array, width, height, rng, accum, n_frames, max_depth, model, n_threads, sun)
        finally:
            set_parity_libm(False)

    def _render(self, cam_array, width, height, rng, accum, n_frames, max_depth, model, n_threads, sun):
        cam = self.camera(cam_array)
        cfg = OrConfig(width, height, max_depth, model, n_threads)
        if sun is not None:
            cfg.sun_light, cfg.sun_dir, cfg.sun_radius = 1, (C.c_float * 3)(*sun[0]), float(sun[1])
        c = OrCounters()
        assert rng.dtype == np.uint32 and rng.flags.c_contiguous and accum.dtype == np.float32 and accum.flags.c_contiguous
        self.l.or_render(C.byref(self.s), C.byref(cam), C.byref(cfg), _p(rng), _p(accum), n_frames, C.byref(c))
        return c.as_dict()

    def ref_glsl_render(self, cam_array, width, height, rng, accum, model=0, sun=None):
        """One dispatch of the reference's own main() (Iterative.comp:214-307, compiled as C++ by oracle/Makefile) over a whole
        frame of this scene, in place on rng and accum like render(). Its texture units and cube map are this library's
        or_sample_texture / or_sample_environment (GL's fixed-function filtering has one statement), everything else on the
        path is the reference's source. Path length is the shader's own (i > 64), i.e. max_depth 66. sun = (direction, radius)
        runs the shader's SUN_LIGHT build."""
        g = C.CDLL(GLSL_SO)
        g.ref_glsl_bind.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
        g.ref_glsl_bind_lights.argtypes = [C.c_void_p, C.c_int, C.c_float]
        g.ref_glsl_bind_materials.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        g.ref_glsl_render_frame.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_int]
        g.ref_glsl_render_frame_sun.argtypes = g.ref_glsl_render_frame.argtypes + [C.c_void_p, C.c_float]
        g.ref_glsl_bind(_p(self.nodes), _p(self.refs), _p(self.tris), self.tris.shape[0])
        g.ref_glsl_bind_lights(_p(self.lights), self.lights.shape[0], float(self.s.total_light_area))
        g.ref_glsl_bind_materials(_p(self.materials), self.materials.shape[0], C.addressof(self.s),
                                  C.cast(self.l.or_sample_texture, C.c_void_p), C.cast(self.l.or_sample_environment, C.c_void_p))
        cam = np.ascontiguousarray(cam_array, dtype=np.float32)
        assert rng.dtype == np.uint32 and rng.flags.c_contiguous and accum.dtype == np.float32 and accum.flags.c_contiguous
        tla = float(self.s.total_light_area)
        if sun is not None:
            d = np.ascontiguousarray(sun[0], dtype=np.float32)
            g.ref_glsl_render_frame_sun(_p(cam), width, height, tla, _p(rng), _p(accum), model, _p(d), float(sun[1]))
        else:
            g.ref_glsl_render_frame(_p(cam), width, height, tla, _p(rng), _p(accum), model)

    def ref_glsl_render_mp(self, cam_array, width, height, rng, accum, n_frames=1, max_depth=66, model=0, n_procs=0):
        """n_frames dispatches of the reference's main() by n_procs forked processes that share the ray counter the way the
        GPU's invocations share the SSBO word (oracle/ref_glsl.cpp ref_glsl_render_frames_mp): the reference's own code on
        all host cores. Same bits as ref_glsl_render, whichever process traces a pixel."""
        g = C.CDLL(GLSL_SO)
        g.ref_glsl_bind.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
        g.ref_glsl_bind_lights.argtypes = [C.c_void_p, C.c_int, C.c_float]
        g.ref_glsl_bind_materials.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        g.ref_glsl_render_frames_mp.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
        g.ref_glsl_bind(_p(self.nodes), _p(self.refs), _p(self.tris), self.tris.shape[0])
        g.ref_glsl_bind_lights(_p(self.lights), self.lights.shape[0], float(self.s.total_light_area))
        g.ref_glsl_bind_materials(_p(self.materials), self.materials.shape[0], C.addressof(self.s),
                                  C.cast(self.l.or_sample_texture, C.c_void_p), C.cast(self.l.or_sample_environment, C.c_void_p))
        cam = np.ascontiguousarray(cam_array, dtype=np.float32)
        assert rng.dtype == np.uint32 and rng.flags.c_contiguous and accum.dtype == np.float32 and accum.flags.c_contiguous
        rc = g.ref_glsl_render_frames_mp(_p(cam), width, height, float(self.s.total_light_area), _p(rng), _p(accum), model,
                                         max_depth, n_frames, n_procs or (os.cpu_count() or 1))
        if rc != 0:
            raise RuntimeError("ref_glsl_render_frames_mp: a worker process failed")

    def light_probe(self, inputs, rng, model=0):
        """or_light_probe: GenerateLightSample on explicit inputs (n x 20 floats) -> (n x 12 floats, RNG words after)."""
        self.l.or_light_probe.argtypes = [C.POINTER(OrScene), C.POINTER(OrConfig), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        a = np.ascontiguousarray(inputs, dtype=np.float32); r = np.ascontiguousarray(rng, dtype=np.uint32).copy()
        cfg = OrConfig(64, 64, 66, model, 1)
        out = np.zeros((a.shape[0], 12), np.float32)
        self.l.or_light_probe(C.byref(self.s), C.byref(cfg), _p(a), _p(r), a.shape[0], _p(out))
        return out, r

    def raygen(self, cam_array, width, height, rng, first, n):
        cam = self.camera(cam_array)
        cfg = OrConfig(width, height, 66, 0, 1)
        out = np.zeros((n, 12), dtype=np.float32)
        self.l.or_raygen(C.byref(cam), C.byref(cfg), float(self.s.total_light_area), _p(rng), first, n, _p(out))
        return out


def hybrid_taus(state4, n):
    st = np.ascontiguousarray(state4, dtype=np.uint32).copy()
    out = np.zeros(n, dtype=np.float32)
    lib().or_hybrid_taus(_p(st), n, _p(out))
    return out, st


def image_plane(aspect, fov, focus, aperture, pos, rot):
    c = OrCamera()
    p = np.asarray(pos, dtype=np.float32); r = np.asarray(rot, dtype=np.float32)
    lib().or_image_plane(aspect, fov, focus, aperture, _p(p), _p(r), C.byref(c))
    return np.array(list(c.lower_left) + list(c.horizontal) + list(c.vertical) + list(c.position) + list(c.u) +
                    list(c.v) + [c.lens_radius], dtype=np.float32)


def tonemap(rgb_sum, num_samples):
    a = np.ascontiguousarray(rgb_sum, dtype=np.float32).reshape(-1, 3)
    out = np.zeros_like(a)
    for i in range(a.shape[0]):
        lib().or_tonemap(_p(a[i:i + 1]), num_samples, _p(out[i:i + 1]))
    return out


# ---------------------------------------------------------------- tinyobjloader (the reference's .obj parser), compiled in place
def equirect_to_cube(rgba, size):
    """LoadEnvironmnet's .hdr branch (Renderer.cpp:262-312) on the CPU: (h, w, 4) float32 -> (6, size, size, 4)."""
    rgba = np.ascontiguousarray(rgba, dtype=np.float32)
    out = np.zeros((6, size, size, 4), dtype=np.float32)
    lib().or_equirect_to_cube(_p(rgba), rgba.shape[1], rgba.shape[0], size, _p(out))
    return out


STBI_SO = os.path.join(_HERE, "_ref", "libref_stbi.so")


def have_stbi():
    return os.path.exists(STBI_SO)


def ref_stbi_loadf(path):
    """stbi_loadf(path, &w, &h, &c, 4) of the stb_image vendored in the reference's SOIL2 -> (h, w, 4) float32."""
    l = C.CDLL(STBI_SO)
    w, h, p = C.c_int(), C.c_int(), C.POINTER(C.c_float)()
    l.ref_stbi_loadf.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p]
    l.ref_stbi_free.argtypes = [C.c_void_p]
    if l.ref_stbi_loadf(path.encode(), C.byref(w), C.byref(h), C.byref(p)) != 0:
        return None
    out = np.ctypeslib.as_array(p, shape=(h.value, w.value, 4)).copy()
    l.ref_stbi_free(p)
    return out


def ref_stbi_load(path):
    """stbi_load(path, &w, &h, &c, 4) (what SOIL_load_image(.., SOIL_LOAD_RGBA) yields) -> (h, w, 4) uint8."""
    l = C.CDLL(STBI_SO)
    w, h, p = C.c_int(), C.c_int(), C.POINTER(C.c_uint8)()
    l.ref_stbi_load.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p]
    l.ref_stbi_free.argtypes = [C.c_void_p]
    if l.ref_stbi_load(path.encode(), C.byref(w), C.byref(h), C.byref(p)) != 0:
        return None
    out = np.ctypeslib.as_array(p, shape=(h.value, w.value, 4)).copy()
    l.ref_stbi_free(p)
    return out


TINYOBJ_SO = os.path.join(_HERE, "_ref", "libref_tinyobj.so")


def have_tinyobj():
    return os.path.exists(TINYOBJ_SO)


def ref_tinyobj_load(path):
    """tinyobj::ObjReader (extern/tinyobjloader, as configured at Scene.cpp:125-136) -> raw arrays: corner_idx (c,3)
    [vertex, normal, texcoord], face_mat (f,), shape_first_face (s+1,), v (nv,3), vn, vt (nvt,2), materials (m,11)
    [Kd, Ks, Ke, Ns, illum], texnames [m]."""
    l = C.CDLL(TINYOBJ_SO)
    I, F = C.POINTER(C.c_int), C.POINTER(C.c_float)
    nc, nf, ns, nv, nvn, nvt, nm = (C.c_int() for _ in range(7))
    ci, fm, sf, v, vn, vt, mf, names = I(), I(), I(), F(), F(), F(), F(), C.c_char_p()
    l.ref_tinyobj_load.argtypes = [C.c_char_p] + [C.c_void_p] * 15
    l.ref_tinyobj_free.argtypes = [C.c_void_p]
    rc = l.ref_tinyobj_load(path.encode(), C.byref(nc), C.byref(ci), C.byref(nf), C.byref(fm), C.byref(ns), C.byref(sf),
                            C.byref(nv), C.byref(v), C.byref(nvn), C.byref(vn), C.byref(nvt), C.byref(vt), C.byref(nm),
                            C.byref(mf), C.byref(names))
    if rc != 0:
        raise RuntimeError(f"tinyobj could not parse {path}")

    def arr(p, shape, dt):
        n = int(np.prod(shape))
        return (np.ctypeslib.as_array(p, shape=(max(n, 1),))[:n].astype(dt).reshape(shape)).copy()
    out = dict(corner_idx=arr(ci, (nc.value, 3), np.int32), face_mat=arr(fm, (nf.value,), np.int32),
               shape_first_face=arr(sf, (ns.value + 1,), np.int32), v=arr(v, (nv.value, 3), np.float32),
               vn=arr(vn, (nvn.value, 3), np.float32), vt=arr(vt, (nvt.value, 2), np.float32),
               materials=arr(mf, (nm.value, 11), np.float32), texnames=names.value.decode().split("\n")[:nm.value])
    for p in (ci, fm, sf, v, vn, vt, mf):
        l.ref_tinyobj_free(p)
    return out
