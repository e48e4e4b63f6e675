/*
 * mi_b200.h -- C ABI of libmi_b200.so: the B200-native volumetric path tracing
 * hot path (volpath / volpathaos forward, prbvolpath / prb_volpathaos adjoint)
 * of the mitsuba3-nasa fork.
 *
 * Every entry point below replaces one reference interface; the citation says
 * which (paths relative to the reference checkout):
 *
 *   mib_scene_create / mib_scene_destroy
 *       the flattened result of mi.load_dict(): PluginManager::create_object
 *       (src/core/plugin.cpp:174-211) + Scene ctor (src/render/scene.cpp).
 *   mib_scene_set_buffer / mib_scene_commit
 *       params[key] = ...; params.update() -> parameters_changed()
 *       (src/python/python/util.py:185-232, src/media/heterogeneous.cpp:176-178,
 *        src/volumes/grid.cpp:306-319).
 *   mib_render
 *       SamplingIntegrator::render (src/render/integrator.cpp:150-397) and
 *       ADIntegrator.render (src/python/python/ad/integrators/common.py:47-111).
 *   mib_render_1
 *       SamplingIntegrator::render_1 (src/render/integrator.cpp:399-743) and
 *       ADIntegrator.render_1 (common.py:113-196).
 *   mib_render_backward / mib_render_1_backward
 *       RBIntegrator.render_backward / render_1_backward (common.py:828-1102).
 *   mib_grad_get / mib_grad_zero
 *       dr.grad(params[key]) after dr.backward().
 *   mib_counters_get
 *       no reference counterpart: exact event counters used for the roofline.
 *
 * Plain pointers and sizes only; no torch / pybind types.  All functions return
 * 0 on success, non-zero on failure with a message in mib_last_error().
 * "_dev" variants take / return device pointers (used by bench.py to time the
 * kernels with inputs already resident in HBM).
 */
#ifndef MI_B200_H
#define MI_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MIB_ABI_VERSION 8

#define MIB_MAX_SHAPES   12
#define MIB_MAX_MEDIA    4
#define MIB_MAX_PHASES   4
#define MIB_MAX_BSDFS    8
#define MIB_MAX_EMITTERS 4
#define MIB_MAX_BUFFERS  24

/* ---- enums (plugin types become tags) ---- */
enum { MIB_SPEC_MONO = 0, MIB_SPEC_RGB = 1, MIB_SPEC_MONO_POLARIZED = 2 };
enum { MIB_INTEG_VOLPATH = 0, MIB_INTEG_VOLPATHAOS = 1,
       MIB_INTEG_PRBVOLPATH = 2, MIB_INTEG_PRBVOLPATHAOS = 3,
       MIB_INTEG_VOLPATHMIS = 4 /* src/integrators/volpathmis.cpp with use_spectral_mis; forward, unpolarized */ };
enum { MIB_SHAPE_CUBE = 0, MIB_SHAPE_RECTANGLE = 1, MIB_SHAPE_SPHERE = 2, MIB_SHAPE_MESH = 3 };
enum { MIB_MESH_HAS_NORMALS = 1, MIB_MESH_HAS_TEXCOORDS = 2, MIB_MESH_HAS_AREA_TABLE = 4 };
enum { MIB_TEXTURE_UNIFORM = 0, MIB_TEXTURE_CHECKERBOARD = 1 };
enum { MIB_BSDF_NULL = 0, MIB_BSDF_DIFFUSE = 1, MIB_BSDF_DIELECTRIC = 2, MIB_BSDF_ROUGHDIELECTRIC = 3, MIB_BSDF_THINDIELECTRIC = 4,
       MIB_BSDF_PLASTIC = 5, MIB_BSDF_ROUGHPLASTIC = 6 };
#define MIB_ROUGH_TRANSMITTANCE_RES 64   /* MI_ROUGH_TRANSMITTANCE_RES, include/mitsuba/render/microfacet.h:22 */
enum { MIB_MICROFACET_BECKMANN = 0, MIB_MICROFACET_GGX = 1 };
enum { MIB_MEDIUM_HOMOGENEOUS = 0, MIB_MEDIUM_HETEROGENEOUS = 1 };
enum { MIB_VOLUME_CONST = 0, MIB_VOLUME_GRID = 1 };
enum { MIB_FILTER_NEAREST = 0, MIB_FILTER_TRILINEAR = 1 };
enum { MIB_WRAP_REPEAT = 0, MIB_WRAP_MIRROR = 1, MIB_WRAP_CLAMP = 2 };
enum { MIB_PHASE_ISOTROPIC = 0, MIB_PHASE_HG = 1, MIB_PHASE_RAYLEIGH = 2,
       MIB_PHASE_TABPHASE = 3, MIB_PHASE_TABPHASE_POLARIZED = 4 };
enum { MIB_EMITTER_CONSTANT = 0, MIB_EMITTER_DIRECTIONAL = 1, MIB_EMITTER_AREA = 2, MIB_EMITTER_ENVMAP = 3, MIB_EMITTER_SPOT = 4 };
#define MIB_MAX_ENV_LEVELS 16
enum { MIB_SENSOR_PERSPECTIVE = 0, MIB_SENSOR_DISTANT = 1, MIB_SENSOR_RADIANCEMETER = 2 };
enum { MIB_RFILTER_BOX = 0, MIB_RFILTER_GAUSSIAN = 1, MIB_RFILTER_TENT = 2 };
enum { MIB_RNG_ARGS_LEFT_TO_RIGHT = 0, MIB_RNG_ARGS_RIGHT_TO_LEFT = 1 };

/* ---- POD scene description.  Matrices are 4x4 row-major float. ---- */
typedef struct mib_volume {
    int32_t type;            /* MIB_VOLUME_* */
    int32_t channels;        /* 1 or 3 */
    float   value[3];        /* constvolume value (src/volumes/const.cpp:61-84) */
    int32_t res[3];          /* grid resolution x,y,z (src/volumes/grid.cpp:156-299) */
    int32_t filter;          /* MIB_FILTER_* */
    int32_t wrap;            /* MIB_WRAP_* */
    int32_t buffer;          /* index of the data buffer [z][y][x][c], -1 if none */
    float   to_local[16];    /* world -> [0,1]^3 (src/render/volume.cpp:13-17) */
    float   bbox_min[3];
    float   bbox_max[3];
    float   max_value;       /* filled by mib_scene_commit (grid.cpp:306-319) unless fixed_max is set */
    int32_t fixed_max;       /* gridvolume `max_value` property (grid.cpp:296-297): max_value is the caller's and is kept */
} mib_volume;

typedef struct mib_phase {
    int32_t type;            /* MIB_PHASE_* */
    float   g;               /* hg (src/phase/hg.cpp:50-62) */
    float   depolarization;  /* rayleigh (src/phase/rayleigh.cpp:41-48) */
    int32_t n_nodes;         /* tabulated: number of cos(theta) nodes */
    int32_t buf_nodes;       /* buffer indices: cos(theta) nodes, m11, m12, m33, m34 */
    int32_t buf_m11;
    int32_t buf_m12;
    int32_t buf_m33;
    int32_t buf_m34;
} mib_phase;

typedef struct mib_medium {
    int32_t type;            /* MIB_MEDIUM_* */
    int32_t phase;           /* index into phases */
    int32_t sample_emitters;
    int32_t has_spectral_extinction;
    float   scale;
    mib_volume sigma_t;
    mib_volume albedo;
    float   max_density;     /* filled by mib_scene_commit (heterogeneous.cpp:163) */
} mib_medium;

typedef struct mib_bsdf {
    int32_t type;            /* MIB_BSDF_* */
    float   reflectance[3];  /* diffuse (src/bsdfs/diffuse.cpp) */
    /* the ocean interface of the fork's atmosphere-ocean scenes: */
    float   eta;             /* dielectric / roughdielectric: int_ior / ext_ior (src/bsdfs/dielectric.cpp:210-222) */
    float   inv_eta;         /* dielectric: ext_ior / int_ior (:223); roughdielectric: rcp(eta) (roughdielectric.cpp:245) */
    float   alpha_u;         /* roughdielectric roughness (src/bsdfs/roughdielectric.cpp:196-218) */
    float   alpha_v;
    int32_t distribution;    /* MIB_MICROFACET_* (roughdielectric.cpp:182-193) */
    int32_t sample_visible;  /* roughdielectric.cpp:195; 0: normals drawn from D cos(theta_m) (microfacet.h:246-293), isotropic roughness only */
    /* diffuse reflectance as a checkerboard texture (src/textures/checkerboard.cpp:66-84; SURVEY 8f rank 3: the background
       plane of the fork's participating_media test scenes): color0 where the two uv cells agree, color1 elsewhere */
    int32_t texture;         /* MIB_TEXTURE_*; UNIFORM: `reflectance` */
    float   color0[3];
    float   color1[3];
    float   to_uv[6];        /* rows 0 and 1 of the 3 x 3 `to_uv` transform (uv' = to_uv * (u, v, 1)) */
    /* plastic / roughplastic (src/bsdfs/plastic.cpp:151-200, roughplastic.cpp:169-253; ABI 8): `reflectance` is the
       diffuse_reflectance, `eta` / `alpha_u` / `distribution` as above; the derived members are filled by the host side
       (mitsuba3_nasa_b200/plastic.py) as the plugins' parameters_changed() does */
    float   specular_reflectance[3];
    int32_t has_specular_reflectance;
    int32_t nonlinear;
    float   inv_eta_2;                  /* 1 / eta^2 */
    float   specular_sampling_weight;   /* s_mean / (d_mean + s_mean) */
    float   internal_reflectance;       /* plastic: fresnel_diffuse_reflectance(1 / eta); roughplastic: m_internal_reflectance */
    float   ext_transmittance[MIB_ROUGH_TRANSMITTANCE_RES];   /* roughplastic: eval_transmittance over mu = linspace(0, 1, 64) */
} mib_bsdf;

typedef struct mib_shape {
    int32_t type;            /* MIB_SHAPE_* */
    int32_t bsdf;            /* index into bsdfs */
    int32_t interior_medium; /* -1: none */
    int32_t exterior_medium; /* -1: none */
    int32_t emitter;         /* area emitter index or -1 */
    float   to_world[16];
    float   to_object[16];
    /* rectangle frame (src/shapes/rectangle.cpp:115-125): dp_du, dp_dv, n */
    float   frame_s[3];
    float   frame_t[3];
    float   frame_n[3];
    float   inv_surface_area;
    /* sphere (src/shapes/sphere.cpp): world-space center / radius */
    float   center[3];
    float   radius;
    /* triangle mesh (`obj`, src/shapes/obj.cpp, include/mitsuba/render/mesh.h), world space, packed into ONE
       float buffer: positions [3 V] | normals [3 V] (HAS_NORMALS) | texcoords [2 V] (HAS_TEXCOORDS) | faces [3 F] (uint32
       bit patterns) | BVH nodes [8 N]: bbox min (3), bbox max (3), then for an inner node (left child, right child | 0x80000000)
       and for a leaf (first entry of the order table, count) -- as uint32 bit patterns | order table [F]: face indices in leaf
       order (uint32 bit patterns) | with HAS_AREA_TABLE (meshes that carry an area emitter): triangle areas [F] | their running
       sum [F] (DiscreteDistribution, distr_1d.h:243-262; `inv_surface_area` = 1 / total).  Built by the host side
       (mitsuba3_nasa_b200/mesh.py); node 0 is the root. */
    int32_t mesh_buffer;     /* buffer slot, -1: not a mesh */
    int32_t mesh_vertices;   /* V */
    int32_t mesh_faces;      /* F */
    int32_t mesh_nodes;      /* N */
    int32_t mesh_flags;      /* MIB_MESH_HAS_* */
} mib_shape;

typedef struct mib_emitter {
    int32_t type;            /* MIB_EMITTER_* */
    float   radiance[3];     /* constant radiance / directional irradiance / area radiance */
    float   direction[3];    /* directional: to_world * (0,0,1) */
    float   bsphere_center[3];
    float   bsphere_radius;  /* after set_scene (constant.cpp:70-85, directional.cpp:98-108) */
    int32_t shape;           /* area emitter: owning shape */
    /* environment map (src/emitters/envmap.cpp), ONE float buffer: texels [H][W][3] (W = file width + 1: the
       last column repeats the first, envmap.cpp:217-224; mono variants hold the luminance in all three channels) followed by the
       levels of the luminance warp (Hierarchical2D<Float, 0>, include/mitsuba/core/distr_2d.h:344-500): level 0 = the
       normalised W x H interpolant, levels 1.. = the 2 x 2-blocked MIP hierarchy of patch integrals, each level at
       env_level_offset[l] (in floats from the start of the buffer) with row pitch env_level_width[l].  Built by the host
       side (mitsuba3_nasa_b200/envmap.py). */
    int32_t env_buffer;      /* buffer slot, -1: not an envmap */
    int32_t env_res[2];      /* W, H */
    float   env_scale;
    float   env_to_world[16];
    float   env_to_local[16];
    int32_t env_levels;      /* number of levels incl. level 0 */
    int32_t env_level_offset[MIB_MAX_ENV_LEVELS];
    int32_t env_level_width[MIB_MAX_ENV_LEVELS];
    /* spot light (src/emitters/spot.cpp:89-151,177-214): `radiance` = intensity, position = translation of env_to_world,
       env_to_local = its inverse; falloff between beam_width and cutoff_angle (radians) */
    float   spot_cutoff_angle;
    float   spot_cos_cutoff;
    float   spot_cos_beam;
    float   spot_inv_transition;
} mib_emitter;

typedef struct mib_sensor {
    int32_t type;            /* MIB_SENSOR_* */
    int32_t medium;          /* medium the sensor sits in, -1 none */
    float   to_world[16];
    float   sample_to_camera[16]; /* perspective (src/sensors/perspective.cpp:173-199) */
    float   near_clip, far_clip;
    float   principal_point_offset[2];
    float   target[3];       /* distant: point target (src/sensors/distant.cpp:165-167) */
    int32_t target_type;     /* 0 none, 1 point */
    float   bsphere_center[3];
    float   bsphere_radius;
    int32_t needs_aperture_sample;
    /* film (src/films/hdrfilm.cpp:245-273, src/render/film.cpp:10-51) */
    int32_t film_width, film_height;
    int32_t crop_offset_x, crop_offset_y, crop_width, crop_height;
    int32_t rfilter;         /* MIB_RFILTER_* */
    float   rfilter_radius;
    float   rfilter_stddev;
    int32_t sample_border;
} mib_sensor;

typedef struct mib_integrator {
    int32_t type;            /* MIB_INTEG_* */
    int32_t max_depth;       /* -1 = unbounded (integrator.cpp:1286-1293) */
    int32_t rr_depth;
    int32_t hide_emitters;
    int32_t rng_arg_order;   /* MIB_RNG_ARGS_*  */
    int32_t double_phase_quirk; /* volpath multiplies phase_weight twice (volpath.cpp:244,251) */
    int32_t majorant_res;    /* 0: reference behaviour (one global majorant per heterogeneous medium, heterogeneous.cpp:163-185);
                              * M > 0: "fast" mode, M^3 supergrid of local majorants + DDA free-flight sampling:
                              * unbiased, but no longer the reference's random sequence -- agreement is statistical only */
    int32_t spectral_mis;    /* volpathmis: `use_spectral_mis` (volpathmis.cpp:93, default 1): n x n weight matrices; 0: one spectrum */
    int32_t unbiased_adjoint; /* prb_volpathaos, extension (property `unbiased`, default 0 = the reference): divide the replayed radiance by
                               * the step's weight as prbvolpath does (prbvolpath.py:192-194) instead of prb_volpathaos.py:214-217's
                               * `I = replace_grad(1, weight)`; the gradient then agrees with finite differences of the reference's
                               * own volpathaos (tests/test_oracle_adjoint.py), which the reference's adjoint does not */
} mib_integrator;

typedef struct mib_scene_desc {
    int32_t abi_version;     /* MIB_ABI_VERSION */
    int32_t spectrum;        /* MIB_SPEC_* */
    int32_t n_shapes, n_media, n_phases, n_bsdfs, n_emitters;
    int32_t environment;     /* index of the constant environment emitter or -1 */
    mib_shape      shapes[MIB_MAX_SHAPES];
    mib_medium     media[MIB_MAX_MEDIA];
    mib_phase      phases[MIB_MAX_PHASES];
    mib_bsdf       bsdfs[MIB_MAX_BSDFS];
    mib_emitter    emitters[MIB_MAX_EMITTERS];
    mib_sensor     sensor;
    mib_integrator integrator;
    uint32_t sampler_base_seed; /* independent sampler 'seed' (sampler.cpp:12-20) */
} mib_scene_desc;

/* Event counters accumulated by the kernels (exact, per render call). */
typedef struct mib_counters {
    uint64_t n_samples;      /* camera paths */
    uint64_t n_lookups;      /* gridvolume lookups (main walk + shadow walks) */
    uint64_t n_real;         /* real medium scatter events */
    uint64_t n_null;         /* null collisions in the main walk */
    uint64_t n_nee_steps;    /* tentative collisions inside shadow walks */
    uint64_t n_surface;      /* surface interactions handled */
    double   kernel_ms;      /* CUDA-event time of the last render kernel(s) */
    uint64_t n_launches;     /* kernels launched by the last call */
} mib_counters;

typedef struct mib_scene mib_scene;

const char *mib_last_error(void);
int  mib_abi_version(void);
size_t mib_sizeof_scene_desc(void);
/* bytes of the device-side scene block uploaded to the kernels' constant bank at every launch (descriptor + buffer
 * pointers + table descriptors); reported by bench.py as part of the end-to-end host->device traffic */
size_t mib_sizeof_device_scene(void);

/* device = CUDA ordinal the scene lives on (one process per GPU). */
int  mib_scene_create(const mib_scene_desc *desc, int device, mib_scene **out);
int  mib_scene_destroy(mib_scene *scene);
/* Upload (copy) a float buffer (grid data, phase tables) into slot idx. */
int  mib_scene_set_buffer(mib_scene *scene, int idx, const float *host, size_t count);
int  mib_scene_get_buffer(mib_scene *scene, int idx, float *host, size_t count);
/* Device pointer of a buffer (for torch-side updates without a host round trip). */
int  mib_scene_buffer_dev(mib_scene *scene, int idx, float **dev, size_t *count);
/* Replace the POD description (scalar parameter updates such as albedo, g, scale). */
int  mib_scene_set_desc(mib_scene *scene, const mib_scene_desc *desc);
/* Recompute derived data: grid maxima, majorants, table CDFs. Must be called
   after set_buffer / set_desc and before rendering. */
int  mib_scene_commit(mib_scene *scene);
int  mib_scene_get_desc(mib_scene *scene, mib_scene_desc *out);

/* Restrict rendering to the global lane range [lane_begin, lane_end) of the
   wavefront (multi-GPU sharding by sample range). (0,0) = all. */
int  mib_scene_set_lane_range(mib_scene *scene, uint64_t lane_begin, uint64_t lane_end);
/* Multi-GPU load balance: instead of one contiguous range, the wavefront is cut into chunks of `chunk_lanes`
 * lanes (a multiple of spp keeps pixels whole), dealt in groups of `world`: in group g rank r renders chunk
 * g * world + ((r - g) mod world), so every rank samples the whole image (the rotation by g keeps a rank from always
 * getting the same column band).  world <= 1 switches striping off; setting stripes clears the lane range. */
int  mib_scene_set_lane_stripes(mib_scene *scene, uint32_t rank, uint32_t world, uint64_t chunk_lanes);

/* Forward render into the raw accumulation film [H][W][C] (C = 4: R,G,B,W for
   rgb; I,Q,U,V for mono_polarized; C = 2: I,W for mono) and, if develop != 0,
   the developed image [H][W][Cout] (Cout = 3 rgb, 1 mono, 4 mono_polarized).
   out_host may be NULL (result stays on the device, see mib_film_dev). */
int  mib_render(mib_scene *scene, uint32_t seed, uint32_t spp, int develop, float *out_host);
int  mib_film_dev(mib_scene *scene, float **raw_dev, size_t *raw_count,
                  float **developed_dev, size_t *developed_count);
int  mib_film_develop(mib_scene *scene);      /* (re)develop after an external reduce */
/* Copies the raw (developed = 0) or developed film to the host. */
int  mib_film_read(mib_scene *scene, int developed, float *host);

/* render_1: sums all samples, returns n_channels doubles normalised by
   W*H*spp (mono: I; mono_polarized: I,Q,U,V). */
int  mib_render_1(mib_scene *scene, uint32_t seed, uint32_t spp, double *out4);
/* Un-normalised sums (for multi-GPU reduction before normalising). */
int  mib_render_1_raw(mib_scene *scene, uint32_t seed, uint32_t spp, double *out4);

/* Adjoint.  grad_in_host: [H][W][Cout] image gradient (render_backward) or 4
   floats (render_1_backward: d loss / d (I,Q,U,V)).  Gradients accumulate
   into per-parameter device buffers until mib_grad_zero. */
int  mib_render_backward(mib_scene *scene, uint32_t seed, uint32_t spp, const float *grad_in_host);
int  mib_render_1_backward(mib_scene *scene, uint32_t seed, uint32_t spp, const float *grad_in4);
/* Gradient of buffer slot idx (same shape as the buffer). */
int  mib_grad_get(mib_scene *scene, int idx, float *host, size_t count);
int  mib_grad_dev(mib_scene *scene, int idx, float **dev, size_t *count);
/* Gradients of scalar medium parameters: out[0..2] = d/d albedo (per channel),
   out[3] = d/d sigma_t (const volume), for medium m. */
int  mib_grad_get_scalar(mib_scene *scene, int medium, double *out4);
/* Device pointer of all scalar gradients, [MIB_MAX_MEDIA][4] doubles (for multi-GPU reductions). */
int  mib_grad_scalar_dev(mib_scene *scene, double **dev, size_t *count);
int  mib_grad_zero(mib_scene *scene);

/* Lorenz-Mie phase-matrix tabulation on the device: MiePhaseFunction::eval_mie (src/phase/mie.cpp:117-215), mie()
   (include/mitsuba/render/mie.h:45-222), mueller::mie_scatter (mueller.h:282-295) averaged over a size distribution
   given by its Gauss-Legendre nodes (SizeDistribution::eval_gauss_all, include/mitsuba/render/sizedistr.h:33-47):
   radius[g], quadrature weight[g], normalised density sdf[g].  n_gauss == 1 with sdf == NULL: monodisperse.
   mu[n_mu] = cos(theta) nodes; outputs are the four distinct Mueller entries per node.  All pointers are host memory. */
int  mib_mie_tabulate(int device, float wavelength, float ior_med_re, float ior_med_im, float ior_sph_re, float ior_sph_im,
                      int nmax, const float *radius, const float *weight, const float *sdf, int n_gauss,
                      const float *mu, int n_mu, float *m11, float *m12, float *m33, float *m34);

int  mib_counters_get(mib_scene *scene, mib_counters *out);
/* which kernel family mib_render / mib_render_1 will launch for this scene: 1 = path pool (k_forward_pool), 0 = one thread per
   path (k_forward); bench.py names the measured kernel with it (no reference counterpart) */
int  mib_scene_forward_kernel(mib_scene *scene, int *is_pool);
/* The CUDA stream the scene launches on (cudaStream_t as void*). */
int  mib_scene_stream(mib_scene *scene, void **stream);

/* Test hooks (no reference counterpart): per-lane sample values and device math. */
int  mib_debug_sample_lanes(mib_scene *scene, uint32_t seed, uint32_t spp, uint64_t lane_begin,
                            uint64_t count, float *out_host /* [count][C] */);
int  mib_debug_math(int op /*0 sincos,1 exp,2 log,3 cbrt,4 asin*/, const float *in_host, float *out_host, size_t n);
/* Fresnel blocks of the ocean interface on the device (csrc/dfresnel.cuh; groundwork for the `dielectric` BSDF):
   fresnel / fresnel_polarized (include/mitsuba/render/fresnel.h:34-71,149-190), mueller::specular_reflection /
   specular_transmission (include/mitsuba/render/mueller.h:198-268), refract (fresnel.h:313-316).
   op 0: in = cos_theta_i [n] -> (F, cos_theta_t, eta_it, eta_ti) [n][4];  1: -> (re a_s, im a_s, re a_p, im a_p, cos_theta_t,
   eta_it, eta_ti) [n][7];  2 / 3: -> row-major 4x4 Mueller matrix [n][16];  4: in = (wi, m) [n][6] -> refracted direction [n][3]
   (zero under total internal reflection). */
int  mib_debug_fresnel(int op, const float *in_host, float eta, float *out_host, size_t n);
/* Microfacet distribution of the rough ocean interface on the device (csrc/dmicrofacet.cuh; groundwork for `roughdielectric`,
   SURVEY 8f rank 2): MicrofacetDistribution eval / smith_g1 / pdf / sample, visible normals (include/mitsuba/render/microfacet.h:
   185-440) and Dr.Jit's erf / erfinv / exp2 (ext/drjit/include/drjit/math.h:756-820,1473-1552).  type 0 = beckmann, 1 = ggx.
   op 0: in = (wi, m) [n][6] -> (eval, smith_g1, pdf) [n][3];  1: in = (wi, sample2) [n][5] -> (m, pdf) [n][4];
   2 / 3 / 4: in = x [n] -> erf(x) / erfinv(x) / exp2(x) [n] (type and alphas ignored). */
int  mib_debug_microfacet(int op, int type, float alpha_u, float alpha_v, const float *in_host, float *out_host, size_t n);
/* fast mode: the majorant supergrid of a heterogeneous medium, [M][M][M] (z, y, x), scale included */
int  mib_debug_supergrid(mib_scene *scene, int medium, float *out_host, size_t count);

/* Measured FP32 roof of a device (no reference counterpart; bench.py's `roofline.peak` for the homogeneous scenes, SURVEY 8d
   "builder measures both with an FMA microbenchmark"): independent FFMA chains on every SM, CUDA-event timed, best of `reps`
   launches, in TFLOP/s with FMA = 2 flops. */
int  mib_debug_fp32_peak(int device, int reps, double *tflops_out);
/* Measured L2 roof of a device (no reference counterpart): all SMs stream one L2-resident 32 MiB buffer with 16-byte loads,
   best of `reps` launches, in GB/s.  bench.py reports the grid look-ups of the heterogeneous scenes against it (they hit the
   L2, so the HBM peak is not the roof they can reach). */
int  mib_debug_l2_peak(int device, int reps, double *gbs_out);

/* Allocation-free variant of mib_debug_sample_lanes.  NO COUNTERPART IN THE REFERENCE: added
   for callers that cannot call cudaMalloc (tight serving loops, harnesses that forbid allocation).
   The per-lane scratch lives in a file-scope `static __device__ __align__(256)` array of MIB_NOALLOC_MAX_FLOATS floats
   (4 MiB per GPU that loads the library, held until the process exits); requests with count * C above that are refused
   with a clear error (C = 2 mono, 4 rgb / polarized), as are count == 0 and lane ranges beyond the wavefront.
   The call performs no cudaMalloc / cudaFree.  It is not capturable into a CUDA graph: like the other render entry points
   it synchronises the scene's stream and reads the event counters back.  Every call of the process shares the array, which
   keeps the last call's lanes until the next call (the part a call uses is cleared first); the library serialises the
   calls of one process, two calls overlapping on one GPU would race. */
#define MIB_NOALLOC_MAX_FLOATS (1u << 20)
int  mib_sample_lanes_noalloc(mib_scene *scene, uint32_t seed, uint32_t spp, uint64_t lane_begin,
                              uint64_t count, float *out_host /* [count][C] */);

#ifdef __cplusplus
}
#endif
#endif /* MI_B200_H */
