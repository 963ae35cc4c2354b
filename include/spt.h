/*
 * spt.h — C ABI of the B200-native wavefront path-tracing core.
 *
 * This is the drop-in boundary for the hot path of PepcyCh/simple-path-tracer.  The reference
 * is a single Rust binary without FFI; the two trait seams this library replaces are
 *
 *   B1  RendererT::render(&self, scene: &Scene, config: &OutputConfig)
 *         src/renderer/mod.rs:16-19   (variant picked by create_renderer, mod.rs:26-38)
 *   B2  PrimitiveT::intersect(&self, ray, inter) -> bool        src/primitive/mod.rs:27
 *       PrimitiveT::intersect_test(&self, ray, t_max) -> bool   src/primitive/mod.rs:25
 *         both invoked on scene.aggregate(), src/core/scene.rs:43-45
 *
 * A Rust host would flatten its `Scene` (src/core/scene.rs:7-12) into an SptSceneDesc once,
 * then call spt_render from a new `Renderer::Wavefront` variant.
 *
 * Conventions: plain C, no exceptions cross the boundary.  Every entry point returns an
 * int32_t status (0 = SPT_OK); the message of the last failure on the calling thread is
 * available from spt_last_error().  All pointers inside descriptors are caller-owned HOST
 * memory that is copied during the call (the library never keeps host pointers), except
 * where a parameter is explicitly named d_* (device pointer).  Handles are opaque and are
 * released by *_destroy.  Calls on one handle are not thread-safe; handles are independent.
 *
 * There is NO CPU fallback: without a CUDA device every compute entry point fails with
 * SPT_ERR_CUDA.
 */
#ifndef SPT_H_
#define SPT_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPT_OK 0
#define SPT_ERR_INVALID 1 /* bad argument / inconsistent descriptor */
#define SPT_ERR_CUDA 2    /* CUDA runtime error (incl. "no device") */
#define SPT_ERR_UNSUPPORTED 3

/* ---- enumerations (values are part of the ABI) ------------------------------------------ */

/* primitive kinds — src/primitive/mod.rs:40-53 (only the in-scope leaves) */
enum { SPT_PRIM_SPHERE = 0, SPT_PRIM_TRIMESH = 1 };

/* material tags — src/material/mod.rs:33-45 (P-NDF variants out of scope) */
enum {
  SPT_MAT_LAMBERT = 0,      /* tex[0]=albedo                                  lambert.rs:26-31  */
  SPT_MAT_PSEUDO = 1,       /* no parameters                                   pseudo.rs:20-24   */
  SPT_MAT_CONDUCTOR = 2,    /* tex[0]=ior tex[1]=ior_k tex[2]=rough_x tex[3]=rough_y  conductor.rs:50-73 */
  SPT_MAT_DIELECTRIC = 3,   /* ior; tex[0]=reflectance tex[1]=transmittance (both ignored, Q11) tex[2..3]=rough */
  SPT_MAT_PLASTIC = 4,      /* ior; tex[0]=albedo tex[2..3]=rough (NOT squared, Q12)  plastic.rs:60-82 */
  SPT_MAT_PBR_METALLIC = 5, /* tex[0]=base_color tex[1]=metallic tex[2..3]=rough  pbr_metallic.rs:67-98 */
  SPT_MAT_PBR_SPECULAR = 6, /* tex[0]=diffuse tex[1]=specular tex[2..3]=rough     pbr_specular.rs:60-88 */
  SPT_MAT_SUBSURFACE = 7    /* ior; tex[0]=albedo tex[1]=ld tex[2..3]=rough        subsurface.rs:65-94  */
};

/* texture tags — src/texture/mod.rs:198-208 */
enum {
  SPT_TEX_SCALAR = 0, /* value[3]                                     scalar.rs:21-51 */
  SPT_TEX_IMAGE = 1,  /* RGBA8 image, mip chain built by the library  image_tex.rs:63-153 */
  SPT_TEX_ADD = 2,    /* child[0] + child[1]                          binary_op.rs:68-73 */
  SPT_TEX_SUB = 3,
  SPT_TEX_MUL = 4,
  SPT_TEX_DIV = 5,
  SPT_TEX_SRGB = 6,          /* srgb->linear of child[0]              srgb_tex.rs:17-69 */
  SPT_TEX_INPUT_MODIFIER = 7 /* tiling/offset/mode/wrap over child[0] input_modifier.rs:34-97 */
};
/* TextureInputMode — src/texture/mod.rs:21-29 ; -1 = "None" in TexInputModifier */
enum { SPT_TEXMODE_SPECIFIED = 0, SPT_TEXMODE_TEXCOORDS = 1, SPT_TEXMODE_POSITION = 2,
       SPT_TEXMODE_NORMAL = 3, SPT_TEXMODE_TANGENT = 4, SPT_TEXMODE_BITANGENT = 5 };
/* TextureInputWrapMode — src/texture/mod.rs:37-43 ; -1 = "None" */
enum { SPT_TEXWRAP_REPEAT = 0, SPT_TEXWRAP_MIRROR_REPEAT = 1, SPT_TEXWRAP_CLAMP = 2,
       SPT_TEXWRAP_MIRROR_CLAMP = 3 };
/* TextureChannel — src/texture/mod.rs:165-171 */
enum { SPT_CHAN_R = 0, SPT_CHAN_G = 1, SPT_CHAN_B = 2, SPT_CHAN_A = 3 };

/* light tags — src/light/mod.rs:28-35 */
enum { SPT_LIGHT_DIRECTIONAL = 0, SPT_LIGHT_POINT = 1, SPT_LIGHT_SPOT = 2, SPT_LIGHT_SHAPE = 3,
       SPT_LIGHT_ENV = 4 };
/* light sampler — src/light_sampler/mod.rs:37-41 */
enum { SPT_LS_UNIFORM = 0, SPT_LS_POWER_IS = 1 };
/* pixel sampler — src/pixel_sampler/mod.rs */
enum { SPT_SAMPLER_RANDOM = 0, SPT_SAMPLER_JITTERED = 1, SPT_SAMPLER_RECURRENCE = 2 };
/* aggregate — src/core/scene_resources.rs:85-103 (informational: results are aggregate-independent) */
enum { SPT_AGG_GROUP = 0, SPT_AGG_BVH = 1 };

/* ---- scene description: flat, index-linked mirror of core/scene.rs:7-12 ----------------- */

typedef struct SptMeshDesc { /* TriMesh — src/primitive/triangle.rs:10-27 */
  int32_t n_verts, n_tris;
  const float* position;  /* 3*n_verts */
  const float* normal;    /* 3*n_verts */
  const float* texcoord;  /* 2*n_verts */
  const float* tangent;   /* 3*n_verts */
  const float* bitangent; /* 3*n_verts */
  const uint32_t* index;  /* 3*n_tris  */
} SptMeshDesc;

typedef struct SptSphereDesc { /* src/primitive/sphere.rs:8-12 */
  float center[3];
  float radius;
} SptSphereDesc;

typedef struct SptInstanceDesc { /* src/primitive/instance.rs:10-16 */
  int32_t prim_kind;  /* SPT_PRIM_* */
  int32_t prim_index; /* into spheres[] or meshes[] */
  int32_t surface;    /* into surfaces[] */
  float xform[12];    /* glam::Affine3A, column-major: x_axis, y_axis, z_axis, translation */
} SptInstanceDesc;

typedef struct SptSurfaceDesc { /* src/core/surface.rs:14-22 */
  int32_t material;
  int32_t normal_map; /* texture or -1 */
  float emissive[3];
  int32_t emissive_map; /* texture or -1 */
  int32_t double_sided;
  int32_t inside_medium; /* medium or -1 */
} SptSurfaceDesc;

typedef struct SptMaterialDesc { /* src/material/ (one file per tag) */
  int32_t tag;
  int32_t tex[4];
  float ior;          /* int_ior / ext_ior */
  int32_t rough_chan; /* SPT_CHAN_*  (pbr_*: roughness_chan) */
  int32_t metal_chan; /* SPT_CHAN_*  (pbr_metallic: metallic_chan) */
} SptMaterialDesc;

typedef struct SptTextureDesc { /* src/texture/ */
  int32_t tag;
  int32_t child[2];
  float value[3];
  float tiling[3], offset[3];
  int32_t mode, wrap;  /* SPT_TEXMODE_* / SPT_TEXWRAP_* or -1 */
  int32_t width, height;
  const uint8_t* rgba; /* width*height*4, row-major, only SPT_TEX_IMAGE */
} SptTextureDesc;

typedef struct SptMediumDesc { /* src/medium/homogeneous.rs:5-9 (sigma_s as LOADED, i.e. after Q13) */
  float sigma_a[3], sigma_s[3];
  float asymmetric;
} SptMediumDesc;

typedef struct SptLightDesc { /* src/light/; listed in the order of the light sampler's Vec */
  int32_t tag;
  float position[3];
  float direction[3]; /* as loaded (DirLight::new normalises it, SpotLight does not) */
  float strength[3];
  float inner_angle, outer_angle; /* radians */
  int32_t instance;               /* SPT_LIGHT_SHAPE: index into instances[] */
} SptLightDesc;

typedef struct SptEnvDesc { /* src/light/environment.rs:9-16 */
  int32_t width, height;
  const float* rgb; /* height rows of width texels, 3 floats each */
  float scale[3];
} SptEnvDesc;

typedef struct SptCameraDesc { /* src/camera/perspective.rs:5-12 */
  float eye[3], forward[3], up[3];
  float fov; /* radians */
} SptCameraDesc;

typedef struct SptSceneDesc {
  int32_t n_meshes;    const SptMeshDesc* meshes;
  int32_t n_spheres;   const SptSphereDesc* spheres;
  int32_t n_instances; const SptInstanceDesc* instances;
  int32_t n_surfaces;  const SptSurfaceDesc* surfaces;
  int32_t n_materials; const SptMaterialDesc* materials;
  int32_t n_textures;  const SptTextureDesc* textures;
  int32_t n_mediums;   const SptMediumDesc* mediums;
  int32_t n_lights;    const SptLightDesc* lights;
  int32_t n_cameras;   const SptCameraDesc* cameras;
  const SptEnvDesc* env;    /* NULL = none */
  int32_t env_light_index;  /* Q8 (scene_resources.rs:165); -1 = none */
  int32_t light_sampler;    /* SPT_LS_* */
  int32_t aggregate;        /* SPT_AGG_* */
} SptSceneDesc;

/* ---- ray queries (seam B2) ---------------------------------------------------------------- */

typedef struct SptRay { /* core/ray.rs:1-27; direction is NOT required to be unit length */
  float origin[3];
  float t_min;
  float direction[3];
  float t_max; /* Intersection::t on entry (f32::MAX = unbounded), core/intersection.rs:21-26 */
} SptRay;

typedef struct SptHit { /* the identity part of core/intersection.rs:6-18 */
  float t;         /* hit distance; t_max of the ray on a miss */
  float u, v;      /* triangle: barycentric weights of vertex 1 and 2 (reference's v,w; triangle.rs:124-147);
                      sphere: texcoords (sphere.rs:138-145) */
  int32_t prim_id; /* triangle index inside its mesh (file order), 0 for a sphere, -1 on a miss */
  int32_t inst_id; /* index into SptSceneDesc::instances, -1 on a miss */
} SptHit;

/* ---- BxDF test hook (BxdfT::{sample,pdf,bxdf,is_delta}, src/bxdf/mod.rs:80-90) -------------- */

typedef struct SptBxdfQuery {
  int32_t material;  /* index into materials[] */
  float texcoord[2]; /* inter.texcoords fed to MaterialT::bxdf_context */
  float wo[3];       /* local frame, z = shading normal */
  float wi[3];       /* direction for pdf()/bxdf() */
  float rand[8];     /* the uniform_1d() draws, consumed in call order by sample() */
} SptBxdfQuery;

typedef struct SptBxdfResult {
  float sample_wi[3];
  float sample_f[3];
  float sample_pdf;
  int32_t sample_lobe; /* 0 diffuse 1 glossy 2 specular  (BxdfLobeType) */
  int32_t sample_dir;  /* 0 reflect 1 transmit           (BxdfDirType)  */
  int32_t n_rand_used;
  float eval_f[3]; /* bxdf(wo, wi) */
  float eval_pdf;  /* pdf(wo, wi)  */
  int32_t is_delta;
  int32_t bxdf_tag; /* which Bxdf variant bxdf_context() built (bxdf/mod.rs:92-102 order) */
} SptBxdfResult;

/* ---- render (seam B1) ------------------------------------------------------------------------- */

#define SPT_RENDER_COUNT_TRAVERSAL 1u /* fill nodes_visited/tris_tested/insts_visited (slower build) */
#define SPT_RENDER_NO_SORT 2u         /* disable material binning of the shade stage (for A/B timing) */
#define SPT_RENDER_TIME_STAGES 4u     /* fill seconds_stage[] from CUDA events around every kernel launch (implies SERIAL) */
#define SPT_RENDER_SERIAL 8u          /* issue every kernel on one stream (default: independent kernels of an iteration overlap) */
#define SPT_RENDER_DEBUG_NORMAL 32u   /* the reference's cargo feature `debug_normal` (pt.rs:113-118): every camera path ends at its first hit and
                                         writes 0.5 * inter.normal + 0.5 (misses: the environment, or black); the cheapest first-hit parity image */
#define SPT_RENDER_ACCUMULATE 16u     /* progressive rendering: keep the film of the previous call on this scene (same width x height)
                                         and add this call's samples to it; the outputs are the mean over all samples so far */

/* Default screen tile = the 8x4 pixel block one warp generates camera samples for.  Measured on C2 at 8 ranks: per-rank
 * render times spread by 8 % with 32x32 tiles (255 tiles per rank), by 1.3 % with 8x4 (8100 per rank). */
#define SPT_DEFAULT_TILE_W 8
#define SPT_DEFAULT_TILE_H 4

typedef struct SptRenderDesc { /* OutputConfig (renderer/mod.rs:9-14) + PathTracer fields (pt.rs:24-28) */
  int32_t camera_index;
  int32_t width, height;
  int32_t max_depth;       /* scenes/pt.json:3 */
  int32_t sampler_kind;    /* SPT_SAMPLER_* */
  int32_t spp;             /* random / recurrence */
  int32_t div_x, div_y;    /* jittered (spp = div_x*div_y) */
  float filter_radius;     /* box filter radius in pixels, filter/boxf.rs, in (0, 3.5].  0.5 = own-pixel mean.  Any other radius reproduces
                              Film::filter_pixel as written (film.rs:68-91, Q14): the UNWEIGHTED colour sum of every sample of the
                              (2R+1)^2 neighbour pixels, R = ceil(radius - 0.5), divided by the number of those samples that lie inside the
                              box — brighter than a mean for every radius but 0.5; a pixel with no sample in the box is inf / NaN.
                              The sample offsets are recomputed from (seed, sampler, pixel, sample index), not stored: progressive passes
                              that end in a radius != 0.5 must cover consecutive sample ranges with the SAME seed and sampler */
  uint64_t seed;           /* RNG key; the stream of a path depends only on (seed, pixel, sample) */
  int32_t tile_w, tile_h;  /* screen tile size for rank interleaving (0 = SPT_DEFAULT_TILE_W x _H); multiples of 8 / 4 */
  int32_t rank, world_size;/* this process renders the tiles it owns; others stay 0 */
  int32_t pool_size;       /* resident path slots, 184 B each (0 = default: 1/6 of this rank's camera samples at first, widened after three iterations to
                              what ends the render in ~10 iterations — at most 1/4 of the samples — all kept within [2^22, 2^25];
                              fewer for renders of less than 2^22 samples, halved if the device is short of memory) */
  uint32_t flags;
  int32_t spp_first, spp_count; /* render only the samples [spp_first, spp_first + spp_count) of every pixel's spp (0, 0 = all).
                                   Sample k of a pixel is the same path whichever call renders it (streams and sampler offsets are
                                   keyed by the sample index), so passes over consecutive ranges with SPT_RENDER_ACCUMULATE add up,
                                   bit for bit, to the frame rendered in one call — also across a spt_film_save / spt_film_load. */
} SptRenderDesc;

typedef struct SptStats {
  uint64_t paths;          /* camera samples traced by this rank */
  uint64_t ext_rays;       /* closest-hit rays (pt.rs:50) */
  uint64_t shadow_rays;    /* any-hit rays actually traced (pt.rs:81,169) */
  uint64_t probe_rays;     /* BSSRDF probes (substrate.rs:290) */
  uint64_t nodes_visited, tris_tested, insts_visited; /* only with SPT_RENDER_COUNT_TRAVERSAL */
  uint64_t nonfinite;      /* contributions dropped because they were NaN/Inf */
  uint64_t iterations;     /* wavefront iterations */
  uint64_t kernel_launches;
  double seconds_total;    /* device time of the whole render (CUDA events) */
  double seconds_stage[6]; /* generate, extend, shade, connect, resolve, other */
  uint64_t shadow_culled;  /* NEE samples whose contribution was exactly zero (light behind the surface, black emitter):
                              the reference still calls intersect_test for them (pt.rs:169); this core skips the
                              any-hit ray because its result is multiplied by zero.  Not counted in shadow_rays. */
  uint64_t trav_profile[10]; /* only with SPT_RENDER_COUNT_TRAVERSAL: lane-rounds with a ray, lanes active in the pop /
                                node / instance / triangle stage, then warp-rounds total and warp-rounds that ran each stage */
} SptStats;

typedef struct SptSceneInfo {
  int64_t tlas_nodes, blas_nodes, triangles, instances;
  int64_t bytes_nodes, bytes_tris, bytes_total;
  double seconds_bvh;     /* acceleration-structure build inside spt_scene_create (all BLASes + TLAS) */
  double seconds_create;  /* whole spt_scene_create: flatten + build + upload */
  int64_t blas_on_device; /* how many BLASes were built by the GPU builder (big meshes) rather than the host SAH builder */
} SptSceneInfo;

typedef struct SptContext SptContext;
typedef struct SptScene SptScene;

const char* spt_last_error(void);
int32_t spt_version(void);

/* One context = one CUDA device (one process per GPU; ranks are joined with spt_comm_create below). */
int32_t spt_context_create(int32_t device_id, SptContext** out);
int32_t spt_context_destroy(SptContext* ctx);

/* Flatten + host wide-BVH build + upload.  Replaces BvhAccel::new (primitive/bvh.rs:23-141),
 * TriMesh::new (triangle.rs:42-55), Instance::new (instance.rs:19-29), EnvLight::new
 * (environment.rs:18-50), PowerIsLightSampler::new (power_is.rs:19-42). */
int32_t spt_scene_create(SptContext* ctx, const SptSceneDesc* desc, SptScene** out);
/* Re-send the already flattened scene arrays host->device (pinned staging; no rebuild). */
int32_t spt_scene_upload(SptScene* scene, int64_t* bytes_copied);
int32_t spt_scene_info(const SptScene* scene, SptSceneInfo* out);
int32_t spt_scene_destroy(SptScene* scene);

/* ≙ PrimitiveT::intersect on scene.aggregate() for n rays (host buffers). */
int32_t spt_trace_batch(SptScene* scene, const SptRay* rays, int64_t n, SptHit* out_hits);
/* ≙ PrimitiveT::intersect_test (ray.t_max is the t_max argument); out[i] = 1 if occluded. */
int32_t spt_occluded_batch(SptScene* scene, const SptRay* rays, int64_t n, uint8_t* out);
/* Same, buffers already resident in device memory; stats may be NULL.  `stream` is a cudaStream_t passed as void*: the batch is
 * enqueued on it and the call returns without synchronising (unless SPT_RENDER_COUNT_TRAVERSAL asks for stats).  stream == NULL
 * means the scene's own internal (non-blocking) stream, and the call then returns only after the batch has finished.
 * Concurrency: batches issued on different caller streams may overlap on the device (each launch has its own fetch cursor, up to
 * 64 launches in flight); calls on ONE scene handle must still come from one host thread at a time, and
 * SPT_RENDER_COUNT_TRAVERSAL batches (shared counters) must not overlap each other or a render. */
int32_t spt_trace_batch_device(SptScene* scene, const SptRay* d_rays, int64_t n, SptHit* d_hits,
                               uint32_t flags, SptStats* stats, void* stream);
int32_t spt_occluded_batch_device(SptScene* scene, const SptRay* d_rays, int64_t n, uint8_t* d_out,
                                  uint32_t flags, SptStats* stats, void* stream);

/* spt_trace_batch without any allocation.  NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs, outputs and
 * results as spt_trace_batch (host buffers), for n <= SPT_NOALLOC_MAX_RAYS rays per call; a larger n returns SPT_ERR_INVALID with a
 * message naming the limit.  The device scratch the batch needs (ray staging, hit staging, the fetch cursor: 3.3 MB) lives in a
 * file-scope device array that every GPU loading the library holds until exit; the call performs no cudaMalloc / cudaFree and calls no
 * allocating library routine.  All calls on a device share that array, which keeps the last call's contents (the call clears what
 * must start at zero): calls that overlap on the GPU would race, so the call synchronises before it returns and must not be issued
 * from two host threads at once. */
#define SPT_NOALLOC_MAX_RAYS 65536
int32_t spt_trace_batch_noalloc(SptScene* scene, const SptRay* rays, int64_t n, SptHit* out_hits);

/* Test hook ≙ BxdfT::{sample,pdf,bxdf,is_delta} with caller-supplied random numbers. */
int32_t spt_bxdf_batch(SptScene* scene, const SptBxdfQuery* queries, int64_t n, SptBxdfResult* out);

/* ≙ RendererT::render.  out_rgb_mean: width*height*3 f32, row 0 = top (film.rs:64-66);
 * out_rgb8 (optional): the 8-bit encode of film.rs:94-99.  Both are HOST buffers. */
int32_t spt_render(SptScene* scene, const SptRenderDesc* desc, float* out_rgb_mean, uint8_t* out_rgb8,
                   SptStats* stats);
/* Same, result left in device memory (width*height*3 f32; pixels of other ranks are 0) so the
 * host can reduce it over NCCL without a round trip. */
int32_t spt_render_device(SptScene* scene, const SptRenderDesc* desc, float* d_rgb_mean, SptStats* stats);

/* ---- multi-GPU (one process per GPU) ---------------------------------------------------------------------------------------------
 * The rank join and the framebuffer exchange live behind this ABI, so a host binds ONE library: rank 0 calls spt_comm_unique_id and
 * hands the SPT_COMM_ID_BYTES bytes to the other ranks by whatever means it has (a file, a socket, MPI, a torch.distributed
 * broadcast), every rank calls spt_comm_create (ncclCommInitRank on its context's device; collective: all ranks must call it), and
 * then spt_render_gather per frame.  NCCL is loaded at run time (dlopen of libnccl.so.2, or of the file SPT_NCCL_LIB names); without it
 * these calls return SPT_ERR_UNSUPPORTED and everything else works.  A process holds ONE libnccl.so.2: a host that also uses another
 * NCCL consumer with its own copy (PyTorch) must load that consumer before the first spt_comm_* call.  The reference has no counterpart: it threads row bands in one process
 * (renderer/pt.rs:243-261). */
#define SPT_COMM_ID_BYTES 128
typedef struct SptComm SptComm;
int32_t spt_comm_unique_id(uint8_t* out_id /* SPT_COMM_ID_BYTES */);
int32_t spt_comm_create(SptContext* ctx, const uint8_t* id /* SPT_COMM_ID_BYTES */, int32_t rank, int32_t world_size, SptComm** out);
int32_t spt_comm_destroy(SptComm* comm);
/* ≙ RendererT::render over all ranks of `comm`: every rank renders the interleaved tiles it owns (desc->rank / world_size are taken
 * from the communicator), then ONLY the owned tiles travel — as the integer film, 3 x int64 per pixel, by grouped ncclSend/ncclRecv
 * over NVLink — to `root`, which unpacks them into its film and reconstructs the frame into d_rgb_mean (DEVICE memory,
 * width*height*3 f32; ignored on the other ranks, may be NULL there).  Because the film is integer the result is bit-identical to the
 * one-GPU frame for every filter radius.  stats (may be NULL) are this rank's; seconds_stage[5] = device time of exchange + resolve.
 * Collective: every rank must call it with the same desc; a rank that fails before the exchange (e.g. out of memory) leaves the others
 * waiting in NCCL, as with any collective. */
int32_t spt_render_gather(SptScene* scene, SptComm* comm, const SptRenderDesc* desc, int32_t root, float* d_rgb_mean, SptStats* stats);

/* Progressive dump / resume: the film is 3 x int64 per pixel (sums in 2^-24 fixed point), row 0 = top.  spt_film_save copies the
 * film of the last render on this scene to the host (n_values = width*height*3); spt_film_load installs one as if `spp_done`
 * samples per pixel had been accumulated, so the next SPT_RENDER_ACCUMULATE call continues from it. */
int32_t spt_film_save(SptScene* scene, int64_t* out_values, int64_t n_values);
int32_t spt_film_load(SptScene* scene, const int64_t* values, int32_t width, int32_t height, int32_t spp_done);

/* Host-only helper (no device needed): the ids of the screen tiles rank `rank` of `world_size` owns for a
 * width x height image cut into tile_w x tile_h tiles (0 = the default).  Tiles are dealt round-robin in hash order, so
 * every rank gets a spatially interleaved share balanced to +-1 tile (replaces the contiguous row bands of
 * renderer/util.rs:6-19).  Writes at most `cap` ids (row-major tile index) and returns the count in *n_out. */
int32_t spt_owned_tiles(int32_t width, int32_t height, int32_t tile_w, int32_t tile_h, int32_t rank, int32_t world_size,
                        uint32_t* out, int64_t cap, int64_t* n_out);

#ifdef __cplusplus
}
#endif
#endif /* SPT_H_ */
