"""ctypes binding of libb200pt.so (include/b200pt.h) plus a thin mirror of the reference's host classes.

`Renderer`, `Scene` and `Camera` keep the names, argument meaning and call order of the reference's
src/core/Renderer.h:12-53, src/core/Scene.h:19-40 and src/math/Camera.h:14-46 so tests read like a
driver of the reference (Program.cpp:69-152). All compute happens in the CUDA library; there is no
fallback: a missing library or a missing GPU raises.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PT_LIB_PATH") or os.path.join(_HERE, "libb200pt.so")   # PT_LIB_PATH: A/B runs of experimental builds

PT_OK = 0
PT_MICROFACET_GGX, PT_MICROFACET_BECKMANN = 0, 1
PT_BVH_REFERENCE, PT_BVH_WIDE, PT_BVH_WIDE4 = 0, 1, 2
PT_PARITY_LIBM = 1
REFERENCE_MAX_DEPTH = 66      # `i > 64`, Iterative.comp:293


class PtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libb200pt error {code}: {msg}")
        self.code = code


class pt_config(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("max_depth", C.c_int32),
                ("microfacet_model", C.c_int32), ("device", C.c_int32), ("bvh_layout", C.c_int32),
                ("n_slices", C.c_int32), ("sun_light", C.c_int32), ("trace_null_shadow", C.c_int32), ("parity_flags", C.c_int32), ("reserved", C.c_int32 * 2)]


class pt_camera(C.Structure):
    _fields_ = [("lower_left", C.c_float * 3), ("horizontal", C.c_float * 3), ("vertical", C.c_float * 3),
                ("position", C.c_float * 3), ("u", C.c_float * 3), ("v", C.c_float * 3), ("lens_radius", C.c_float)]


class pt_texture(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("rgba", C.POINTER(C.c_float))]


class pt_counters(C.Structure):
    _fields_ = [("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("paths", C.c_uint64),
                ("frames", C.c_uint64), ("kernel_launches", C.c_uint64), ("null_shadow_rays", C.c_uint64),
                ("closest_pair_tests", C.c_uint64), ("closest_tri_tests", C.c_uint64), ("shadow_pair_tests", C.c_uint64),
                ("shadow_tri_tests", C.c_uint64), ("counted_closest_rays", C.c_uint64), ("counted_shadow_rays", C.c_uint64)]


class pt_bvh_stats(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("n_refs", C.c_int32), ("n_leaves", C.c_int32), ("max_depth", C.c_int32),
                ("sah_cost", C.c_float), ("build_ms", C.c_float), ("total_ms", C.c_float)]


class pt_scene_txt(C.Structure):
    _fields_ = [("scene_path", C.c_char * 1024), ("environment", C.c_char * 1024),
                ("camera_position", C.c_float * 3), ("camera_rotation", C.c_float * 3)]


# every symbol include/b200pt.h declares: (name, restype, argtypes)
_P = C.c_void_p
_F = C.POINTER(C.c_float)
SYMBOLS = [
    ("pt_abi_version", C.c_int, []),
    ("pt_last_error", C.c_char_p, []),
    ("pt_create", C.c_int, [C.POINTER(pt_config), C.POINTER(_P)]),
    ("pt_destroy", None, [_P]),
    ("pt_set_triangles", C.c_int, [_P, _P, C.c_int64]),
    ("pt_set_materials", C.c_int, [_P, _P, C.c_int32, _P, C.c_int32]),
    ("pt_set_lights", C.c_int, [_P, _P, C.c_int32, C.c_float]),
    ("pt_set_environment_color", C.c_int, [_P, _F]),
    ("pt_set_sun", C.c_int, [_P, _F, C.c_float]),
    ("pt_set_environment_cubemap", C.c_int, [_P, _P, C.c_int32]),
    ("pt_set_environment_equirect", C.c_int, [_P, _P, C.c_int32, C.c_int32, C.c_int32]),
    ("pt_environment_size", C.c_int32, [_P]),
    ("pt_download_environment", C.c_int, [_P, _P]),
    ("pt_load_environment", C.c_int, [_P, C.c_char_p]),
    ("pt_checkpoint_save", C.c_int, [_P, C.c_char_p]),
    ("pt_checkpoint_load", C.c_int, [_P, C.c_char_p]),
    ("pt_image_save_png", C.c_int, [C.c_char_p, C.c_int32, C.c_int32, _P]),
    ("pt_image_load_hdr", C.c_int, [C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(_F)]),
    ("pt_set_bvh", C.c_int, [_P, _P, C.c_int32, _P, C.c_int32]),
    ("pt_build_bvh", C.c_int, [_P, C.POINTER(pt_bvh_stats)]),
    ("pt_build_bvh_device", C.c_int, [_P, C.POINTER(pt_bvh_stats)]),
    ("pt_build_bvh_device_ploc", C.c_int, [_P, C.c_int32, C.POINTER(pt_bvh_stats)]),
    ("pt_build_bvh_device_split", C.c_int, [_P, C.c_int32, C.c_int32, C.POINTER(pt_bvh_stats)]),
    ("pt_build_bvh_device_sah", C.c_int, [_P, C.POINTER(pt_bvh_stats)]),
    ("pt_optimize_bvh", C.c_int, [_P, C.c_int32, C.c_float, C.POINTER(pt_bvh_stats)]),
    ("pt_bvh_optimize_host", C.c_int, [_P, C.c_int32, _P, C.c_int32, C.c_int32, C.c_float, _F, _F]),
    ("pt_bvh_build_host", C.c_int, [_P, C.c_int64, C.c_int32, C.POINTER(_F), C.POINTER(C.c_int32), C.POINTER(C.POINTER(C.c_int32)), C.POINTER(C.c_int32), C.POINTER(pt_bvh_stats)]),
    ("pt_free", None, [_P]),
    ("pt_bvh_collapse_host", C.c_int, [_P, C.c_int32, _P, C.c_int32, _P, C.c_int64, C.c_int32, C.POINTER(C.POINTER(C.c_uint8)),
                                       C.POINTER(C.c_int64), C.POINTER(C.c_int32), C.POINTER(C.POINTER(C.c_int32)),
                                       C.POINTER(C.c_int64), C.POINTER(C.c_int32)]),
    ("pt_scene_file_load_obj", C.c_int, [C.c_char_p, C.POINTER(_P)]),
    ("pt_scene_file_free", None, [_P]),
    ("pt_scene_file_num_triangles", C.c_int64, [_P]),
    ("pt_scene_file_vertices", _F, [_P]),
    ("pt_scene_file_material_ids", C.POINTER(C.c_uint32), [_P]),
    ("pt_scene_file_num_materials", C.c_int32, [_P]),
    ("pt_scene_file_materials", _F, [_P]),
    ("pt_scene_file_texture_path", C.c_char_p, [_P, C.c_int32]),
    ("pt_scene_file_num_shapes", C.c_int32, [_P]),
    ("pt_scene_file_warnings", C.c_char_p, [_P]),
    ("pt_image_load", C.c_int, [C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(_F)]),
    ("pt_scene_txt_read", C.c_int, [C.c_char_p, _P]),
    ("pt_parse_environment", C.c_int, [C.c_char_p, _F]),
    ("pt_load_scene", C.c_int, [_P, C.c_char_p, C.c_char_p, C.POINTER(pt_bvh_stats)]),
    ("pt_bvh_sah_cost", C.c_float, [_P, C.c_int32, _P, C.c_int32]),
    ("pt_get_bvh", C.c_int, [_P, _P, C.POINTER(C.c_int32), _P, C.POINTER(C.c_int32)]),
    ("pt_assemble_triangles", C.c_int, [_P, _P, C.c_int64, _P]),
    ("pt_build_light_table", C.c_int, [_P, C.c_int64, _P, C.c_int32, _P, C.c_int32, _F]),
    ("pt_presubtract_edges", None, [_P, C.c_int64]),
    ("pt_camera_image_plane", None, [C.c_float, C.c_float, C.c_float, C.c_float, _F, _F, C.POINTER(pt_camera)]),
    ("pt_set_camera", C.c_int, [_P, C.POINTER(pt_camera)]),
    ("pt_set_rng", C.c_int, [_P, _P]),
    ("pt_seed_rng", C.c_int, [_P, C.c_uint64]),
    ("pt_get_rng", C.c_int, [_P, _P]),
    ("pt_render_frames", C.c_int, [_P, C.c_int32]),
    ("pt_reset_samples", C.c_int, [_P]),
    ("pt_num_samples", C.c_uint32, [_P]),
    ("pt_synchronize", C.c_int, [_P]),
    ("pt_download_accum", C.c_int, [_P, _P]),
    ("pt_accum_device_ptr", _P, [_P]),
    ("pt_set_accum_device_ptr", C.c_int, [_P, _P]),
    ("pt_resolve", C.c_int, [_P, C.c_uint32, _P, _P]),
    ("pt_present_async", C.c_int, [_P, C.c_uint32, _P]),
    ("pt_save_screenshot", C.c_int, [_P, C.c_char_p]),
    ("pt_trace_closest", C.c_int, [_P, _P, C.c_int64, _P]),
    ("pt_trace_any", C.c_int, [_P, _P, C.c_int64, _P]),
    ("pt_trace_closest_device", C.c_int, [_P, _P, C.c_int64, _P]),
    ("pt_trace_any_device", C.c_int, [_P, _P, C.c_int64, _P]),
    ("pt_trace_closest_device_noalloc", C.c_int, [_P, _P, C.c_int64, _P]),
    ("pt_noalloc_max_rays", C.c_int64, []),
    ("pt_raygen_probe", C.c_int, [_P, C.c_int64, C.c_int64, _P]),
    ("pt_trace_primary_async", C.c_int, [_P, _P]),
    ("pt_parity_libm_probe", C.c_int, [_P, C.c_int32, _P, _P, C.c_int64, _P]),
    ("pt_get_counters", C.c_int, [_P, C.POINTER(pt_counters)]),
    ("pt_set_visit_counting", C.c_int, [_P, C.c_int]),
    ("pt_last_device_ms", C.c_float, [_P]),
    ("pt_set_stage_timing", C.c_int, [_P, C.c_int]),
    ("pt_get_stage_ms", C.c_int, [_P, _F]),
    ("pt_stream", _P, [_P]),
    ("pt_comm_unique_id", C.c_int, [_P]),
    ("pt_comm_init", C.c_int, [_P, _P, C.c_int32, C.c_int32]),
    ("pt_comm_rank", C.c_int32, [_P]),
    ("pt_comm_world", C.c_int32, [_P]),
    ("pt_comm_nccl_version", C.c_int, []),
    ("pt_rank_seed", C.c_uint64, [C.c_uint64, C.c_int32]),
    ("pt_split_samples", None, [C.c_uint32, C.c_int32, C.c_int32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    ("pt_comm_seed_rng", C.c_int, [_P, C.c_uint64]),
    ("pt_comm_broadcast_bvh", C.c_int, [_P, C.c_int32]),
    ("pt_render_batch", C.c_int, [_P, C.c_int32]),
    ("pt_comm_reset", C.c_int, [_P]),
    ("pt_comm_num_samples", C.c_uint32, [_P]),
    ("pt_comm_synchronize", C.c_int, [_P]),
    ("pt_comm_download_total", C.c_int, [_P, _P]),
    ("pt_comm_download_batch", C.c_int, [_P, _P]),
    ("pt_comm_total_device_ptr", _P, [_P]),
    ("pt_comm_last_reduce_ms", C.c_float, [_P]),
    ("pt_comm_last_batch_ms", C.c_float, [_P]),
    ("pt_comm_present_async", C.c_int, [_P, _P]),
    ("pt_create_multi", C.c_int, [C.POINTER(pt_config), _P, C.c_int32, C.POINTER(_P)]),
    ("pt_group_destroy", None, [_P]),
    ("pt_group_size", C.c_int32, [_P]),
    ("pt_group_context", _P, [_P, C.c_int32]),
    ("pt_group_share_scene", C.c_int, [_P]),
    ("pt_group_set_camera", C.c_int, [_P, C.POINTER(pt_camera)]),
    ("pt_group_seed_rng", C.c_int, [_P, C.c_uint64]),
    ("pt_group_render_batch", C.c_int, [_P, C.c_int32]),
    ("pt_group_reset_samples", C.c_int, [_P]),
    ("pt_group_num_samples", C.c_uint32, [_P]),
    ("pt_group_synchronize", C.c_int, [_P]),
    ("pt_group_download_total", C.c_int, [_P, _P]),
    ("pt_group_present_async", C.c_int, [_P, _P]),
    ("pt_group_save_screenshot", C.c_int, [_P, C.c_char_p]),
]

_lib = None


def load_library():
    """Load libb200pt.so; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: run `python -m gpupathtracer_b200.build` (there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, res, args in SYMBOLS:
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _f3(v):
    return (C.c_float * 3)(*[float(x) for x in v])


def _check(rc):
    if rc != PT_OK:
        raise PtError(rc, load_library().pt_last_error().decode())


# ------------------------------------------------------------------------------- host-side scene assembly
def assemble_triangles(verts, mat_ids):
    """Scene.cpp:263-292 -> (n,20) float32 with ABSOLUTE positions."""
    lib = load_library()
    verts = np.ascontiguousarray(verts, dtype=np.float32)
    mat_ids = np.ascontiguousarray(mat_ids, dtype=np.uint32)
    n = verts.shape[0]
    out = np.zeros((n, 20), dtype=np.float32)
    rc = lib.pt_assemble_triangles(_ptr(verts), _ptr(mat_ids), n, _ptr(out))
    if rc != PT_OK:
        raise PtError(rc, "pt_assemble_triangles")
    return out


def build_light_table(tris_abs, materials):
    """Scene.cpp:296-331 -> ((L,2) float32 {cumArea, triIndex bits}, totalLightArea)."""
    lib = load_library()
    n = tris_abs.shape[0]
    total = C.c_float(0)
    mats = np.ascontiguousarray(materials, dtype=np.float32)
    cnt = lib.pt_build_light_table(_ptr(tris_abs), n, _ptr(mats), mats.shape[0], None, 0, C.byref(total))
    if cnt < 0:
        raise PtError(cnt, "pt_build_light_table")
    out = np.zeros((max(cnt, 1), 2), dtype=np.float32)
    cnt = lib.pt_build_light_table(_ptr(tris_abs), n, _ptr(mats), mats.shape[0], _ptr(out), cnt, C.byref(total))
    return out[:cnt], float(total.value)


def build_bvh_host(tris_pre, n_threads=0):
    """This library's SBVH builder on the host (no device needed) -> (nodes (n,8), refs, stats dict)."""
    lib = load_library()
    t = np.ascontiguousarray(tris_pre, dtype=np.float32)
    nodes_p, refs_p = _F(), C.POINTER(C.c_int32)()
    nn, nr = C.c_int32(), C.c_int32()
    st = pt_bvh_stats()
    rc = lib.pt_bvh_build_host(_ptr(t), t.shape[0], n_threads, C.byref(nodes_p), C.byref(nn), C.byref(refs_p), C.byref(nr), C.byref(st))
    if rc != PT_OK:
        raise PtError(rc, "pt_bvh_build_host")
    nodes = np.ctypeslib.as_array(nodes_p, shape=(nn.value, 8)).copy()
    refs = np.ctypeslib.as_array(refs_p, shape=(nr.value,)).copy()
    lib.pt_free(nodes_p); lib.pt_free(refs_p)
    return nodes, refs, {k: getattr(st, k) for k, _ in pt_bvh_stats._fields_}


def collapse_bvh_host(nodes, refs, tris_pre, width=8):
    """Reference-form tree -> compressed wide layout (host only). Returns (node_bytes (n, node_size) uint8, tri_ids, depth)."""
    lib = load_library()
    nodes = np.ascontiguousarray(nodes, dtype=np.float32); refs = np.ascontiguousarray(refs, dtype=np.int32)
    t = np.ascontiguousarray(tris_pre, dtype=np.float32)
    nb, ids = C.POINTER(C.c_uint8)(), C.POINTER(C.c_int32)()
    n_wide, n_ids, size, depth = C.c_int64(), C.c_int64(), C.c_int32(), C.c_int32()
    rc = lib.pt_bvh_collapse_host(_ptr(nodes), nodes.shape[0], _ptr(refs), refs.shape[0], _ptr(t), t.shape[0], width,
                                  C.byref(nb), C.byref(n_wide), C.byref(size), C.byref(ids), C.byref(n_ids), C.byref(depth))
    if rc != PT_OK:
        raise PtError(rc, "pt_bvh_collapse_host")
    node_bytes = np.ctypeslib.as_array(nb, shape=(n_wide.value, size.value)).copy()
    tri_ids = np.ctypeslib.as_array(ids, shape=(max(n_ids.value, 1),))[:n_ids.value].copy()
    lib.pt_free(nb); lib.pt_free(ids)
    return node_bytes, tri_ids, depth.value


def load_obj(path):
    """LoadOBJ (Scene.cpp:120-231) through libb200pt's own parser -> dict(verts (n,3,8), mat_ids (n,), materials (m,12),
    texture_paths [m], n_shapes, warnings)."""
    lib = load_library()
    h = C.c_void_p()
    _check(lib.pt_scene_file_load_obj(path.encode(), C.byref(h)))
    try:
        n = lib.pt_scene_file_num_triangles(h)
        m = lib.pt_scene_file_num_materials(h)
        verts = np.ctypeslib.as_array(lib.pt_scene_file_vertices(h), shape=(n, 3, 8)).copy()
        mat_ids = np.ctypeslib.as_array(lib.pt_scene_file_material_ids(h), shape=(n,)).copy()
        mats = np.ctypeslib.as_array(lib.pt_scene_file_materials(h), shape=(m, 12)).copy()
        tex = [lib.pt_scene_file_texture_path(h, i).decode() for i in range(m)]
        return dict(verts=verts, mat_ids=mat_ids, materials=mats, texture_paths=tex,
                    n_shapes=int(lib.pt_scene_file_num_shapes(h)), warnings=lib.pt_scene_file_warnings(h).decode())
    finally:
        lib.pt_scene_file_free(h)


def load_image(path):
    """map_Kd decoder -> (h, w, 4) float32."""
    lib = load_library()
    w, h, p = C.c_int32(), C.c_int32(), _F()
    _check(lib.pt_image_load(path.encode(), C.byref(w), C.byref(h), C.byref(p)))
    out = np.ctypeslib.as_array(p, shape=(h.value, w.value, 4)).copy()
    lib.pt_free(p)
    return out


def save_png(path, rgb8):
    """(h, w, 3) uint8, row 0 on top -> 8-bit RGB PNG (the format of Renderer::SaveScreenshot)."""
    rgb8 = np.ascontiguousarray(rgb8, dtype=np.uint8)
    assert rgb8.ndim == 3 and rgb8.shape[2] == 3
    _check(load_library().pt_image_save_png(path.encode(), rgb8.shape[1], rgb8.shape[0], _ptr(rgb8)))


def load_hdr(path):
    """Radiance .hdr as stbi_loadf(path, .., 4) returns it (Renderer.cpp:265) -> (h, w, 4) float32."""
    lib = load_library()
    w, h, p = C.c_int32(), C.c_int32(), _F()
    _check(lib.pt_image_load_hdr(path.encode(), C.byref(w), C.byref(h), C.byref(p)))
    out = np.ctypeslib.as_array(p, shape=(h.value, w.value, 4)).copy()
    lib.pt_free(p)
    return out


def scene_data_from_obj(path, camera=None, env_color=(1.0, 1.0, 1.0)):
    """The file the reference would load (Scene::LoadScene) as a `scenes.SceneData`, textures decoded and bound the way
    pt_load_scene does it -- so the oracle and the per-array upload path can be fed from an .obj too."""
    from .scenes import SceneData
    d = load_obj(path)
    mats = d["materials"].copy()
    textures = []
    for m, tp in enumerate(d["texture_paths"]):
        if not tp:
            continue
        try:
            img = load_image(tp)
        except PtError:
            continue
        textures.append(img)
        mats.view(np.int32)[m, 3] = len(textures) - 1
    return SceneData(os.path.basename(path), d["verts"], d["mat_ids"], mats, tuple(env_color), camera or {}, textures)


def read_scene_txt(path):
    """scene.txt (Program.cpp:70-84) -> (scene_path, environment, camera position, camera rotation)."""
    st = pt_scene_txt()
    _check(load_library().pt_scene_txt_read(path.encode(), C.byref(st)))
    return st.scene_path.decode(), st.environment.decode(), tuple(st.camera_position), tuple(st.camera_rotation)


def parse_environment(args):
    rgb = (C.c_float * 3)()
    _check(load_library().pt_parse_environment(args.encode(), rgb))
    return tuple(rgb)


def presubtract_edges(tris_abs):
    """Scene.cpp:334-337 (returns a copy)."""
    t = np.ascontiguousarray(tris_abs, dtype=np.float32).copy()
    load_library().pt_presubtract_edges(_ptr(t), t.shape[0])
    return t


class Camera:
    """math/Camera.h:14-46. rotation = (yaw, pitch, roll-ignored) as SetRotation (Camera.cpp:33-36)."""

    def __init__(self, aspect, fieldofview, focus, aperture):
        self.aspect, self.fov, self.focus, self.aperture = float(aspect), float(fieldofview), float(focus), float(aperture)
        self.position = (0.0, 0.0, 0.0)
        self.rotation = (0.0, 0.0, 0.0)
        self.plane = pt_camera()

    def SetPosition(self, p):
        self.position = tuple(float(x) for x in p)

    def SetRotation(self, r):
        self.rotation = tuple(float(x) for x in r)

    def GenerateImagePlane(self):
        load_library().pt_camera_image_plane(self.aspect, self.fov, self.focus, self.aperture, _f3(self.position),
                                             _f3(self.rotation), C.byref(self.plane))
        return self.plane

    def as_array(self):
        p = self.plane
        return np.array(list(p.lower_left) + list(p.horizontal) + list(p.vertical) + list(p.position) + list(p.u) +
                        list(p.v) + [p.lens_radius], dtype=np.float32)


class Scene:
    """Host-side result of Scene::LoadScene (Scene.cpp:233-357) for a procedural SceneData."""

    def __init__(self, data):
        self.data = data
        self.tris_abs = assemble_triangles(data.verts, data.mat_ids)
        self.materials = np.ascontiguousarray(data.materials, dtype=np.float32)
        self.lights, self.total_light_area = build_light_table(self.tris_abs, self.materials)
        self.tris = presubtract_edges(self.tris_abs)
        self.env_color = tuple(data.env_color)
        self.env_cubemap = None       # (6, S, S, 4) float32: +X -X +Y -Y +Z -Z faces, replaces env_color when set
        self.textures = [np.ascontiguousarray(t, dtype=np.float32) for t in data.textures]
        self.nodes = None
        self.refs = None

    def set_bvh(self, nodes, refs):
        self.nodes = np.ascontiguousarray(nodes, dtype=np.float32)
        self.refs = np.ascontiguousarray(refs, dtype=np.int32)

    def bounds(self):
        p = self.tris_abs[:, 0:9].reshape(-1, 3)
        return p.min(axis=0), p.max(axis=0)


class Renderer:
    """src/core/Renderer.h:12-53, headless. Initialize(Window*, scenePath, env) becomes
    Initialize(width, height, scene) where `scene` is a `Scene`; everything else keeps its name."""

    def __init__(self):
        self.ctx = None
        self.lib = load_library()
        self.width = self.height = 0

    # -- Renderer::Initialize, Renderer.cpp:448-636
    def Initialize(self, width, height, scene, max_depth=REFERENCE_MAX_DEPTH, microfacet_model=PT_MICROFACET_GGX,
                   device=0, rng_state=None, seed=None, bvh="given", n_slices=0, bvh_layout=PT_BVH_REFERENCE, sun_light=False,
                   trace_null_shadow=False, parity_libm=False):
        cfg = pt_config(width=width, height=height, max_depth=max_depth, microfacet_model=microfacet_model,
                        device=device, bvh_layout=bvh_layout, n_slices=n_slices, sun_light=1 if sun_light else 0,
                        trace_null_shadow=1 if trace_null_shadow else 0, parity_flags=PT_PARITY_LIBM if parity_libm else 0)
        ctx = C.c_void_p()
        _check(self.lib.pt_create(C.byref(cfg), C.byref(ctx)))
        self.ctx = ctx
        self.width, self.height = width, height
        self.scene = scene
        _check(self.lib.pt_set_triangles(ctx, _ptr(scene.tris), scene.tris.shape[0]))
        texs = (pt_texture * max(1, len(scene.textures)))()
        for i, t in enumerate(scene.textures):
            texs[i].width, texs[i].height = t.shape[1], t.shape[0]
            texs[i].rgba = t.ctypes.data_as(C.POINTER(C.c_float))
        _check(self.lib.pt_set_materials(ctx, _ptr(scene.materials), scene.materials.shape[0],
                                         C.cast(texs, C.c_void_p), len(scene.textures)))
        _check(self.lib.pt_set_lights(ctx, _ptr(scene.lights), scene.lights.shape[0], scene.total_light_area))
        _check(self.lib.pt_set_environment_color(ctx, _f3(scene.env_color)))
        if getattr(scene, "env_cubemap", None) is not None:
            self.SetEnvironmentCubemap(scene.env_cubemap)
        if bvh in ("device", "ploc", "split", "split_lbvh", "sah"):
            stats = pt_bvh_stats()
            if bvh in ("split", "split_lbvh"):      # device build with spatial splits: PLOC (or the Morton hierarchy) over pre-split references
                _check(self.lib.pt_build_bvh_device_split(ctx, int(os.environ.get("PT_PLOC_RADIUS", "16")) if bvh == "split" else 0,
                                                          int(os.environ.get("PT_SPLIT_GRID", "64")), C.byref(stats)))
            elif bvh == "sah":                       # top-down binned SAH on the device
                _check(self.lib.pt_build_bvh_device_sah(ctx, C.byref(stats)))
            elif bvh == "ploc":
                _check(self.lib.pt_build_bvh_device_ploc(ctx, int(os.environ.get("PT_PLOC_RADIUS", "16")), C.byref(stats)))
            else:
                _check(self.lib.pt_build_bvh_device(ctx, C.byref(stats)))
            self.bvh_stats = stats
            nodes, refs = self.GetBVH()            # reads the serialised tree back from the device
            scene.set_bvh(nodes, refs)
            stats.sah_cost = self.lib.pt_bvh_sah_cost(_ptr(scene.nodes), scene.nodes.shape[0], _ptr(scene.refs), scene.refs.shape[0])
        elif bvh == "none":
            self.bvh_stats = pt_bvh_stats()         # the tree arrives later (CommBroadcastBVH / pt_set_bvh)
        elif bvh == "build" or scene.nodes is None:
            stats = pt_bvh_stats()
            _check(self.lib.pt_build_bvh(ctx, C.byref(stats)))
            self.bvh_stats = stats
            nodes, refs = self.GetBVH()
            scene.set_bvh(nodes, refs)
        else:
            _check(self.lib.pt_set_bvh(ctx, _ptr(scene.nodes), scene.nodes.shape[0], _ptr(scene.refs), scene.refs.shape[0]))
        if rng_state is not None:
            self.SetRNG(rng_state)
        elif seed is not None:
            _check(self.lib.pt_seed_rng(ctx, seed))
        return self

    def InitializeFromFile(self, width, height, scene_path, env="GENERATE COLOR WHITE", max_depth=REFERENCE_MAX_DEPTH,
                           microfacet_model=PT_MICROFACET_GGX, device=0, seed=1, bvh_layout=PT_BVH_REFERENCE, n_slices=0):
        """Renderer::Initialize(Window*, scenePath, env) (Renderer.cpp:448-636) with the window reduced to its size."""
        cfg = pt_config(width=width, height=height, max_depth=max_depth, microfacet_model=microfacet_model,
                        device=device, bvh_layout=bvh_layout, n_slices=n_slices)
        ctx = C.c_void_p()
        _check(self.lib.pt_create(C.byref(cfg), C.byref(ctx)))
        self.ctx = ctx
        self.width, self.height = width, height
        stats = pt_bvh_stats()
        _check(self.lib.pt_load_scene(ctx, scene_path.encode(), env.encode(), C.byref(stats)))
        self.bvh_stats = stats
        _check(self.lib.pt_seed_rng(ctx, seed))
        return self

    # -- checkpoint / resume (no counterpart in the reference)
    def SaveCheckpoint(self, path):
        _check(self.lib.pt_checkpoint_save(self.ctx, path.encode()))

    def LoadCheckpoint(self, path):
        _check(self.lib.pt_checkpoint_load(self.ctx, path.encode()))

    def SetSun(self, direction, radius):
        """uniforms sunDir / sunRadius (Iterative.comp:89-90); only used with Initialize(.., sun_light=True)."""
        _check(self.lib.pt_set_sun(self.ctx, _f3(direction), float(radius)))

    # -- LoadEnvironmnet, Renderer.cpp:236-323
    def SetEnvironmentCubemap(self, faces):
        faces = np.ascontiguousarray(faces, dtype=np.float32)
        assert faces.ndim == 4 and faces.shape[0] == 6 and faces.shape[1] == faces.shape[2] and faces.shape[3] == 4
        _check(self.lib.pt_set_environment_cubemap(self.ctx, _ptr(faces), faces.shape[1]))

    def SetEnvironmentEquirect(self, rgba, cube_size=1024):
        rgba = np.ascontiguousarray(rgba, dtype=np.float32)
        assert rgba.ndim == 3 and rgba.shape[2] == 4
        _check(self.lib.pt_set_environment_equirect(self.ctx, _ptr(rgba), rgba.shape[1], rgba.shape[0], cube_size))

    def LoadEnvironment(self, args):
        _check(self.lib.pt_load_environment(self.ctx, args.encode()))

    def GetEnvironment(self):
        n = self.lib.pt_environment_size(self.ctx)
        if n == 0:
            return None
        out = np.zeros((6, n, n, 4), dtype=np.float32)
        _check(self.lib.pt_download_environment(self.ctx, _ptr(out)))
        return out

    def GetBVH(self):
        nn, nr = C.c_int32(), C.c_int32()
        _check(self.lib.pt_get_bvh(self.ctx, None, C.byref(nn), None, C.byref(nr)))
        nodes = np.zeros((nn.value, 8), dtype=np.float32)
        refs = np.zeros(nr.value, dtype=np.int32)
        _check(self.lib.pt_get_bvh(self.ctx, _ptr(nodes), C.byref(nn), _ptr(refs), C.byref(nr)))
        return nodes, refs

    def OptimizeBVH(self, iterations=4, batch_fraction=0.1):
        """pt_optimize_bvh: reinsertion optimisation of the installed tree (host pass); the scene object follows the new tree."""
        stats = pt_bvh_stats()
        _check(self.lib.pt_optimize_bvh(self.ctx, int(iterations), float(batch_fraction), C.byref(stats)))
        if getattr(self, "scene", None) is not None:
            self.scene.set_bvh(*self.GetBVH())
        self.opt_stats = stats
        return stats

    def SetRNG(self, state):
        state = np.ascontiguousarray(state, dtype=np.uint32)
        assert state.size == self.width * self.height * 4
        _check(self.lib.pt_set_rng(self.ctx, _ptr(state)))

    def GetRNG(self):
        out = np.zeros((self.width * self.height, 4), dtype=np.uint32)
        _check(self.lib.pt_get_rng(self.ctx, _ptr(out)))
        return out

    # -- Renderer::RenderFrame(const Camera&), Renderer.cpp:651-662: +1 sample per pixel
    def RenderFrame(self, camera, n_frames=1):
        _check(self.lib.pt_set_camera(self.ctx, C.byref(camera.plane)))
        _check(self.lib.pt_render_frames(self.ctx, n_frames))

    def Synchronize(self):
        _check(self.lib.pt_synchronize(self.ctx))

    def ResetSamples(self):
        _check(self.lib.pt_reset_samples(self.ctx))

    def GetNumSamples(self):
        return int(self.lib.pt_num_samples(self.ctx))

    # -- Renderer::Present, Renderer.cpp:681-685 -> headless resolve
    def Present(self, as_u8=False):
        n = self.width * self.height
        if as_u8:
            out = np.zeros((self.height, self.width, 3), dtype=np.uint8)
            _check(self.lib.pt_resolve(self.ctx, self.GetNumSamples(), None, _ptr(out)))
        else:
            out = np.zeros((self.height, self.width, 3), dtype=np.float32)
            _check(self.lib.pt_resolve(self.ctx, self.GetNumSamples(), _ptr(out), None))
        return out

    def SaveScreenshot(self, filename):
        _check(self.lib.pt_save_screenshot(self.ctx, filename.encode()))

    def DownloadAccum(self):
        out = np.zeros((self.height, self.width, 4), dtype=np.float32)
        _check(self.lib.pt_download_accum(self.ctx, _ptr(out)))
        return out

    def TraceClosest(self, rays):
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        hits = np.zeros((rays.shape[0], 4), dtype=np.float32)
        _check(self.lib.pt_trace_closest(self.ctx, _ptr(rays), rays.shape[0], _ptr(hits)))
        return hits

    def TraceAny(self, rays):
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        flags = np.zeros(rays.shape[0], dtype=np.uint8)
        _check(self.lib.pt_trace_any(self.ctx, _ptr(rays), rays.shape[0], _ptr(flags)))
        return flags

    def TraceClosestDevice(self, d_rays, n, d_hits):
        """Device-resident batch (raw device pointers); asynchronous on the library's stream."""
        _check(self.lib.pt_trace_closest_device(self.ctx, C.c_void_p(d_rays), n, C.c_void_p(d_hits)))

    def TraceClosestDeviceNoAlloc(self, d_rays, n, d_hits):
        """pt_trace_closest_device_noalloc: the binned trace with its scratch in a fixed __device__ array (no allocation)."""
        _check(self.lib.pt_trace_closest_device_noalloc(self.ctx, C.c_void_p(d_rays), n, C.c_void_p(d_hits)))

    def TraceAnyDevice(self, d_rays, n, d_flags):
        _check(self.lib.pt_trace_any_device(self.ctx, C.c_void_p(d_rays), n, C.c_void_p(d_flags)))

    def TraceClosestHost(self, h_rays, n, h_hits):
        """Host batch given as raw (ideally pinned) host pointers: H2D, trace, D2H, synchronous."""
        _check(self.lib.pt_trace_closest(self.ctx, C.c_void_p(h_rays), n, C.c_void_p(h_hits)))

    def TraceAnyHost(self, h_rays, n, h_flags):
        _check(self.lib.pt_trace_any(self.ctx, C.c_void_p(h_rays), n, C.c_void_p(h_flags)))

    def TracePrimaryAsync(self, host_ptr):
        """pt_trace_primary_async into (preferably pinned) host memory of width * height * 16 bytes; complete after Synchronize()."""
        _check(self.lib.pt_trace_primary_async(self.ctx, C.c_void_p(host_ptr)))

    def TracePrimary(self):
        """One primary ray per pixel as RenderFrame's first bounce traces it -> (height * width, 4) hit records in pixel order."""
        out = np.zeros((self.width * self.height, 4), dtype=np.float32)
        _check(self.lib.pt_trace_primary_async(self.ctx, _ptr(out)))
        _check(self.lib.pt_synchronize(self.ctx))
        return out

    def RaygenProbe(self, first, n):
        out = np.zeros((n, 12), dtype=np.float32)
        _check(self.lib.pt_raygen_probe(self.ctx, first, n, _ptr(out)))
        return out

    def ParityLibmProbe(self, fn, x, y=None):
        """pt_parity_libm.h on the device: fn in ('sin', 'cos', 'log', 'exp', 'pow')."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        yy = np.ascontiguousarray(y, dtype=np.float32) if y is not None else None
        out = np.zeros_like(x)
        _check(self.lib.pt_parity_libm_probe(self.ctx, ["sin", "cos", "log", "exp", "pow"].index(fn), _ptr(x),
                                             _ptr(yy) if yy is not None else None, x.size, _ptr(out)))
        return out

    def Counters(self):
        c = pt_counters()
        _check(self.lib.pt_get_counters(self.ctx, C.byref(c)))
        return {k: int(getattr(c, k)) for k, _ in pt_counters._fields_}

    def SetVisitCounting(self, on=True):
        """Instrumented traversal kernels: Counters() then carries pair / triangle tests of the rays traced meanwhile."""
        _check(self.lib.pt_set_visit_counting(self.ctx, 1 if on else 0))

    def LastDeviceMs(self):
        return float(self.lib.pt_last_device_ms(self.ctx))

    def SetStageTiming(self, on=True):
        _check(self.lib.pt_set_stage_timing(self.ctx, 1 if on else 0))

    def StageMs(self):
        out = (C.c_float * 5)()
        _check(self.lib.pt_get_stage_ms(self.ctx, out))
        return dict(zip(["raygen", "extend", "shade", "shadow", "tally"], [float(x) for x in out]))

    def ResolveInto(self, host_ptr_u8, num_samples=None):
        """Present into caller-owned (ideally pinned) host memory: height*width*3 bytes."""
        _check(self.lib.pt_resolve(self.ctx, num_samples or self.GetNumSamples(), None, C.c_void_p(host_ptr_u8)))

    def PresentAsync(self, host_ptr_u8, num_samples=None):
        """Present() whose D2H copy overlaps the next frame (pt_present_async); complete after Synchronize()."""
        _check(self.lib.pt_present_async(self.ctx, num_samples or self.GetNumSamples(), C.c_void_p(host_ptr_u8)))

    # -- several GPUs, one process per GPU (pt_comm_*): samples split across ranks, one NCCL reduce per batch
    @staticmethod
    def CommUniqueId():
        buf = (C.c_uint8 * 128)()
        _check(load_library().pt_comm_unique_id(buf))
        return bytes(buf)

    def CommInit(self, unique_id, rank, world):
        buf = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        _check(self.lib.pt_comm_init(self.ctx, buf, rank, world))
        self.rank, self.world = rank, world

    def CommBroadcastBVH(self, root=0):
        _check(self.lib.pt_comm_broadcast_bvh(self.ctx, root))

    def CommSeedRNG(self, seed):
        _check(self.lib.pt_comm_seed_rng(self.ctx, seed))

    def RenderBatch(self, camera, n_frames):
        """n_frames on this rank + the batch's reduce to rank 0 (collective: every rank calls it)."""
        _check(self.lib.pt_set_camera(self.ctx, C.byref(camera.plane)))
        _check(self.lib.pt_render_batch(self.ctx, n_frames))

    def CommReset(self):
        _check(self.lib.pt_comm_reset(self.ctx))

    def CommNumSamples(self):
        return int(self.lib.pt_comm_num_samples(self.ctx))

    def CommSynchronize(self):
        _check(self.lib.pt_comm_synchronize(self.ctx))

    def CommDownloadTotal(self):
        out = np.zeros((self.height, self.width, 4), dtype=np.float32)
        _check(self.lib.pt_comm_download_total(self.ctx, _ptr(out)))
        return out

    def CommDownloadBatch(self):
        out = np.zeros((self.height, self.width, 4), dtype=np.float32)
        _check(self.lib.pt_comm_download_batch(self.ctx, _ptr(out)))
        return out

    def CommPresentAsync(self, host_ptr_u8):
        _check(self.lib.pt_comm_present_async(self.ctx, C.c_void_p(host_ptr_u8)))

    def CommLastBatchMs(self):
        return float(self.lib.pt_comm_last_batch_ms(self.ctx))

    def CommLastReduceMs(self):
        return float(self.lib.pt_comm_last_reduce_ms(self.ctx))

    def AccumDevicePtr(self):
        return int(self.lib.pt_accum_device_ptr(self.ctx))

    def CleanUp(self):
        if self.ctx is not None:
            self.lib.pt_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.CleanUp()
        except Exception:
            pass


class _Borrowed(Renderer):
    """A Renderer view of a context owned by a MultiRenderer's group (never destroys it)."""

    def __init__(self, ctx, width, height):
        super().__init__()
        self.ctx, self.width, self.height = ctx, width, height

    def CleanUp(self):
        self.ctx = None


class MultiRenderer:
    """One process, one host thread, n GPUs (pt_create_multi): the Renderer interface with the samples of every frame split
    across the GPUs. RenderFrame(camera, n) renders n frames on EVERY GPU (n x n_gpus samples per pixel) and reduces them to
    GPU 0, whose running total Present / DownloadAccum / GetNumSamples report."""

    def __init__(self):
        self.lib = load_library()
        self.group = None

    def Initialize(self, width, height, scene, n_gpus, device_ids=None, seed=1, bvh="given", **kw):
        cfg = pt_config(width=width, height=height, max_depth=kw.get("max_depth", REFERENCE_MAX_DEPTH),
                        microfacet_model=kw.get("microfacet_model", PT_MICROFACET_GGX), device=0,
                        bvh_layout=kw.get("bvh_layout", PT_BVH_REFERENCE), n_slices=kw.get("n_slices", 0),
                        sun_light=1 if kw.get("sun_light") else 0, trace_null_shadow=1 if kw.get("trace_null_shadow") else 0,
                        parity_flags=PT_PARITY_LIBM if kw.get("parity_libm") else 0)
        ids = (C.c_int32 * n_gpus)(*(device_ids or range(n_gpus)))
        g = C.c_void_p()
        _check(self.lib.pt_create_multi(C.byref(cfg), ids, n_gpus, C.byref(g)))
        self.group, self.width, self.height, self.n_gpus = g, width, height, n_gpus
        self.scene = scene
        ctx0 = C.c_void_p(self.lib.pt_group_context(g, 0))
        r0 = self.rank(0)
        _check(self.lib.pt_set_triangles(ctx0, _ptr(scene.tris), scene.tris.shape[0]))
        texs = (pt_texture * max(1, len(scene.textures)))()
        for i, t in enumerate(scene.textures):
            texs[i].width, texs[i].height = t.shape[1], t.shape[0]
            texs[i].rgba = t.ctypes.data_as(C.POINTER(C.c_float))
        _check(self.lib.pt_set_materials(ctx0, _ptr(scene.materials), scene.materials.shape[0], C.cast(texs, C.c_void_p), len(scene.textures)))
        _check(self.lib.pt_set_lights(ctx0, _ptr(scene.lights), scene.lights.shape[0], scene.total_light_area))
        _check(self.lib.pt_set_environment_color(ctx0, _f3(scene.env_color)))
        if getattr(scene, "env_cubemap", None) is not None:
            r0.SetEnvironmentCubemap(scene.env_cubemap)
        stats = pt_bvh_stats()
        if bvh == "build" or scene.nodes is None:
            _check(self.lib.pt_build_bvh(ctx0, C.byref(stats)))
            scene.set_bvh(*r0.GetBVH())
        else:
            _check(self.lib.pt_set_bvh(ctx0, _ptr(scene.nodes), scene.nodes.shape[0], _ptr(scene.refs), scene.refs.shape[0]))
        self.bvh_stats = stats
        _check(self.lib.pt_group_share_scene(g))          # the other GPUs get copies; the tree is never rebuilt
        _check(self.lib.pt_group_seed_rng(g, seed))
        return self

    def rank(self, i):
        return _Borrowed(C.c_void_p(self.lib.pt_group_context(self.group, i)), self.width, self.height)

    def RenderFrame(self, camera, n_frames=1):
        _check(self.lib.pt_group_set_camera(self.group, C.byref(camera.plane)))
        _check(self.lib.pt_group_render_batch(self.group, n_frames))

    def ResetSamples(self):
        _check(self.lib.pt_group_reset_samples(self.group))

    def GetNumSamples(self):
        return int(self.lib.pt_group_num_samples(self.group))

    def Synchronize(self):
        _check(self.lib.pt_group_synchronize(self.group))

    def DownloadAccum(self):
        out = np.zeros((self.height, self.width, 4), dtype=np.float32)
        _check(self.lib.pt_group_download_total(self.group, _ptr(out)))
        return out

    def PresentAsync(self, host_ptr_u8):
        _check(self.lib.pt_group_present_async(self.group, C.c_void_p(host_ptr_u8)))

    def SaveScreenshot(self, filename):
        _check(self.lib.pt_group_save_screenshot(self.group, filename.encode()))

    def CleanUp(self):
        if self.group is not None:
            self.lib.pt_group_destroy(self.group)
            self.group = None

    def __del__(self):
        try:
            self.CleanUp()
        except Exception:
            pass


def make_rng_table(n_pixels, seed=0):
    """Stand-in for Renderer.cpp:545-553 (std::default_random_engine + uniform_int_distribution(129, UINT32_MAX));
    the table is an INPUT shared by the CUDA path and the oracle."""
    rng = np.random.default_rng(seed)
    return rng.integers(129, 2 ** 32, size=(n_pixels, 4), dtype=np.uint64).astype(np.uint32)
