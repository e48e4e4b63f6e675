"""ctypes mirror of include/mi_b200.h (the C ABI of libmi_b200.so).

Keep in sync with the header; `check_abi(lib)` compares sizeof(mib_scene_desc)
and the ABI version at load time so a drift fails loudly.
"""
import ctypes as C

MIB_ABI_VERSION = 8
MIB_MAX_SHAPES, MIB_MAX_MEDIA, MIB_MAX_PHASES = 12, 4, 4
MIB_MAX_BSDFS, MIB_MAX_EMITTERS, MIB_MAX_BUFFERS = 8, 4, 24

SPEC_MONO, SPEC_RGB, SPEC_MONO_POLARIZED = 0, 1, 2
INTEG_VOLPATH, INTEG_VOLPATHAOS, INTEG_PRBVOLPATH, INTEG_PRBVOLPATHAOS, INTEG_VOLPATHMIS = 0, 1, 2, 3, 4
SHAPE_CUBE, SHAPE_RECTANGLE, SHAPE_SPHERE, SHAPE_MESH = 0, 1, 2, 3
MESH_HAS_NORMALS, MESH_HAS_TEXCOORDS, MESH_HAS_AREA_TABLE = 1, 2, 4
TEXTURE_UNIFORM, TEXTURE_CHECKERBOARD = 0, 1
BSDF_NULL, BSDF_DIFFUSE, BSDF_DIELECTRIC, BSDF_ROUGHDIELECTRIC, BSDF_THINDIELECTRIC, BSDF_PLASTIC, BSDF_ROUGHPLASTIC = 0, 1, 2, 3, 4, 5, 6
MIB_ROUGH_TRANSMITTANCE_RES = 64
MICROFACET_BECKMANN, MICROFACET_GGX = 0, 1
MEDIUM_HOMOGENEOUS, MEDIUM_HETEROGENEOUS = 0, 1
VOLUME_CONST, VOLUME_GRID = 0, 1
FILTER_NEAREST, FILTER_TRILINEAR = 0, 1
WRAP_REPEAT, WRAP_MIRROR, WRAP_CLAMP = 0, 1, 2
PHASE_ISOTROPIC, PHASE_HG, PHASE_RAYLEIGH, PHASE_TABPHASE, PHASE_TABPHASE_POLARIZED = 0, 1, 2, 3, 4
EMITTER_CONSTANT, EMITTER_DIRECTIONAL, EMITTER_AREA, EMITTER_ENVMAP, EMITTER_SPOT = 0, 1, 2, 3, 4
MIB_MAX_ENV_LEVELS = 16
SENSOR_PERSPECTIVE, SENSOR_DISTANT, SENSOR_RADIANCEMETER = 0, 1, 2
RFILTER_BOX, RFILTER_GAUSSIAN, RFILTER_TENT = 0, 1, 2
RNG_ARGS_LEFT_TO_RIGHT, RNG_ARGS_RIGHT_TO_LEFT = 0, 1

i32, f32 = C.c_int32, C.c_float


class Volume(C.Structure):
    _fields_ = [("type", i32), ("channels", i32), ("value", f32 * 3), ("res", i32 * 3),
                ("filter", i32), ("wrap", i32), ("buffer", i32), ("to_local", f32 * 16),
                ("bbox_min", f32 * 3), ("bbox_max", f32 * 3), ("max_value", f32), ("fixed_max", i32)]


class Phase(C.Structure):
    _fields_ = [("type", i32), ("g", f32), ("depolarization", f32), ("n_nodes", i32),
                ("buf_nodes", i32), ("buf_m11", i32), ("buf_m12", i32), ("buf_m33", i32), ("buf_m34", i32)]


class Medium(C.Structure):
    _fields_ = [("type", i32), ("phase", i32), ("sample_emitters", i32), ("has_spectral_extinction", i32),
                ("scale", f32), ("sigma_t", Volume), ("albedo", Volume), ("max_density", f32)]


class Bsdf(C.Structure):
    _fields_ = [("type", i32), ("reflectance", f32 * 3), ("eta", f32), ("inv_eta", f32), ("alpha_u", f32), ("alpha_v", f32),
                ("distribution", i32), ("sample_visible", i32),
                ("texture", i32), ("color0", f32 * 3), ("color1", f32 * 3), ("to_uv", f32 * 6),
                ("specular_reflectance", f32 * 3), ("has_specular_reflectance", i32), ("nonlinear", i32), ("inv_eta_2", f32),
                ("specular_sampling_weight", f32), ("internal_reflectance", f32), ("ext_transmittance", f32 * MIB_ROUGH_TRANSMITTANCE_RES)]


class Shape(C.Structure):
    _fields_ = [("type", i32), ("bsdf", i32), ("interior_medium", i32), ("exterior_medium", i32),
                ("emitter", i32), ("to_world", f32 * 16), ("to_object", f32 * 16),
                ("frame_s", f32 * 3), ("frame_t", f32 * 3), ("frame_n", f32 * 3), ("inv_surface_area", f32),
                ("center", f32 * 3), ("radius", f32),
                ("mesh_buffer", i32), ("mesh_vertices", i32), ("mesh_faces", i32), ("mesh_nodes", i32), ("mesh_flags", i32)]


class Emitter(C.Structure):
    _fields_ = [("type", i32), ("radiance", f32 * 3), ("direction", f32 * 3),
                ("bsphere_center", f32 * 3), ("bsphere_radius", f32), ("shape", i32),
                ("env_buffer", i32), ("env_res", i32 * 2), ("env_scale", f32), ("env_to_world", f32 * 16), ("env_to_local", f32 * 16),
                ("env_levels", i32), ("env_level_offset", i32 * 16), ("env_level_width", i32 * 16),
                ("spot_cutoff_angle", f32), ("spot_cos_cutoff", f32), ("spot_cos_beam", f32), ("spot_inv_transition", f32)]


class Sensor(C.Structure):
    _fields_ = [("type", i32), ("medium", i32), ("to_world", f32 * 16), ("sample_to_camera", f32 * 16),
                ("near_clip", f32), ("far_clip", f32), ("principal_point_offset", f32 * 2),
                ("target", f32 * 3), ("target_type", i32), ("bsphere_center", f32 * 3), ("bsphere_radius", f32),
                ("needs_aperture_sample", i32),
                ("film_width", i32), ("film_height", i32), ("crop_offset_x", i32), ("crop_offset_y", i32),
                ("crop_width", i32), ("crop_height", i32), ("rfilter", i32), ("rfilter_radius", f32),
                ("rfilter_stddev", f32), ("sample_border", i32)]


class Integrator(C.Structure):
    _fields_ = [("type", i32), ("max_depth", i32), ("rr_depth", i32), ("hide_emitters", i32),
                ("rng_arg_order", i32), ("double_phase_quirk", i32), ("majorant_res", i32), ("spectral_mis", i32), ("unbiased_adjoint", i32)]


class SceneDesc(C.Structure):
    _fields_ = [("abi_version", i32), ("spectrum", i32),
                ("n_shapes", i32), ("n_media", i32), ("n_phases", i32), ("n_bsdfs", i32), ("n_emitters", i32),
                ("environment", i32),
                ("shapes", Shape * MIB_MAX_SHAPES), ("media", Medium * MIB_MAX_MEDIA),
                ("phases", Phase * MIB_MAX_PHASES), ("bsdfs", Bsdf * MIB_MAX_BSDFS),
                ("emitters", Emitter * MIB_MAX_EMITTERS), ("sensor", Sensor), ("integrator", Integrator),
                ("sampler_base_seed", C.c_uint32)]


class Counters(C.Structure):
    _fields_ = [("n_samples", C.c_uint64), ("n_lookups", C.c_uint64), ("n_real", C.c_uint64),
                ("n_null", C.c_uint64), ("n_nee_steps", C.c_uint64), ("n_surface", C.c_uint64),
                ("kernel_ms", C.c_double), ("n_launches", C.c_uint64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


# Every symbol include/mi_b200.h declares (tests check the library exports all of them).
EXPORTS = [
    "mib_last_error", "mib_abi_version", "mib_sizeof_scene_desc", "mib_sizeof_device_scene",
    "mib_scene_create", "mib_scene_destroy", "mib_scene_set_buffer", "mib_scene_get_buffer",
    "mib_scene_buffer_dev", "mib_scene_set_desc", "mib_scene_commit", "mib_scene_get_desc",
    "mib_scene_set_lane_range", "mib_scene_set_lane_stripes", "mib_render", "mib_film_dev", "mib_film_develop", "mib_film_read",
    "mib_render_1", "mib_render_1_raw", "mib_render_backward", "mib_render_1_backward",
    "mib_grad_get", "mib_grad_dev", "mib_grad_get_scalar", "mib_grad_scalar_dev", "mib_grad_zero",
    "mib_mie_tabulate", "mib_counters_get", "mib_scene_forward_kernel", "mib_scene_stream",
    "mib_debug_sample_lanes", "mib_debug_math", "mib_debug_fresnel", "mib_debug_microfacet", "mib_debug_supergrid", "mib_debug_fp32_peak", "mib_debug_l2_peak",
    "mib_sample_lanes_noalloc",
]


def check_abi(lib):
    lib.mib_abi_version.restype = C.c_int
    lib.mib_sizeof_scene_desc.restype = C.c_size_t
    lib.mib_sizeof_device_scene.restype = C.c_size_t
    if lib.mib_abi_version() != MIB_ABI_VERSION:
        raise RuntimeError("libmi_b200.so ABI version %d != python mirror %d" % (lib.mib_abi_version(), MIB_ABI_VERSION))
    if lib.mib_sizeof_scene_desc() != C.sizeof(SceneDesc):
        raise RuntimeError("sizeof(mib_scene_desc) mismatch: C %d vs ctypes %d" %
                           (lib.mib_sizeof_scene_desc(), C.sizeof(SceneDesc)))
