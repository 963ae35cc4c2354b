// spt_api.cu — the extern "C" entry points of include/spt.h: context, scene upload, ray batches, render loop.
#include <cstdio>
#include <chrono>
#include <cstring>
#include <cstdlib>
#include <string>
#include <vector>
#include <algorithm>
#include <new>
#include <cmath>
#include <mutex>
#include <dlfcn.h>
#include <nccl.h> // types and enums only: the functions are bound at run time (nccl_load)
#include "../../include/spt.h"
#include "flatten.h"
#include "launch.h"

#define SPT_BATCH_CURSORS 64

namespace {

thread_local std::string g_err;
int fail(int code, const std::string& m) { g_err = m; return code; }
#define CU(call)                                                                                        \
  do {                                                                                                  \
    cudaError_t e_ = (call);                                                                            \
    if (e_ != cudaSuccess) return fail(SPT_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
  } while (0)

struct DevBuf { // one device allocation + the pinned staging copy used by spt_scene_upload
  void* d = nullptr; void* h = nullptr; size_t bytes = 0;
  const void* src = nullptr; // big arrays are not staged: they are copied straight from the HostScene vector the scene keeps
  bool in_arena = false;     // d points into SptScene::hot_arena (not freed by itself)
};

} // namespace

struct SptContext { int device = 0; int sm_count = 148; };

struct SptScene {
  SptContext* ctx = nullptr;
  spt::HostScene hs;
  DScene ds{};
  std::vector<DevBuf> bufs;
  std::vector<DCamera> cameras;
  // the arrays traversal reads (nodes, tris, tlas_inst, inst) live in ONE allocation, so that one L2 access-policy window can cover them
  char* hot_arena = nullptr; size_t hot_bytes = 0, hot_used = 0; bool l2_window = false;
  // render-time buffers, grown on demand and reused across spt_render calls
  void* pool_mem = nullptr; size_t pool_bytes = 0; int pool_n = 0;
  long long* film = nullptr; size_t film_px = 0; // film_px = allocated pixels
  int film_w = 0, film_h = 0, film_spp = 0;       // what the film currently holds: frame size and samples per pixel accumulated
  int film_k0 = 0; bool film_contiguous = true;   // ... and which: the samples [film_k0, film_k0 + film_spp) of every pixel, if the passes were consecutive
  DevCounters* ctr = nullptr;
  uint32_t* batch_next = nullptr; uint32_t batch_seq = 0; // ring of SPT_BATCH_CURSORS ray-fetch cursors: one per batch launch in flight
  void* stage_rays = nullptr; void* stage_out = nullptr; size_t stage_rays_bytes = 0, stage_out_bytes = 0; // device staging of the host-buffer entry points (grown on demand, kept)
  uint32_t* tiles = nullptr; size_t tiles_cap = 0, n_tiles = 0; int tile_key[6] = {0, 0, 0, 0, 0, 0};
  float* out_mean = nullptr; uint8_t* out_rgb8 = nullptr; size_t out_px = 0;
  RenderParams last_P{};                             // parameters of the last render (what the film holds): the resolve of spt_render_gather needs them
  long long* gather_buf = nullptr; size_t gather_bytes = 0;   // spt_render_gather: tile-packed film of this rank (root: of every rank)
  uint32_t* all_tiles = nullptr; size_t all_tiles_cap = 0; int all_tiles_key[6] = {0, 0, 0, 0, 0, 0}; // root: every rank's tile list, rank-major, padded to the same length
  DevCounters* h_ctr = nullptr; // pinned
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[2] = {nullptr, nullptr};
  cudaEvent_t sev[4] = {nullptr, nullptr, nullptr, nullptr}; // per-stage timing (SPT_RENDER_TIME_STAGES)
  int occ_ext = 0, occ_con = 0, occ_bat = 0;
  cudaStream_t aux[3] = {nullptr, nullptr, nullptr};         // [0],[1]: shade kernels of other material classes, [2]: connect
  cudaEvent_t ev_fork = nullptr, ev_shaded = nullptr, ev_con = nullptr, ev_join[2] = {nullptr, nullptr}; // resident blocks per SM of the persistent kernels (occupancy API)
};

namespace {

template <typename T>
int upload(SptScene* s, const std::vector<T>& v, const T** dptr, bool kept = true, bool hot = false) { // kept: v lives as long as the scene (a HostScene member)
  DevBuf b; b.bytes = std::max<size_t>(sizeof(T), v.size() * sizeof(T));
  if (hot && s->hot_arena && s->hot_used + b.bytes <= s->hot_bytes) { b.d = s->hot_arena + s->hot_used; b.in_arena = true; s->hot_used += (b.bytes + 255) & ~(size_t)255; }
  else CU(cudaMalloc(&b.d, b.bytes));
  const size_t used = v.size() * sizeof(T);
  s->bufs.push_back(b); // owned by the scene from here on: spt_scene_destroy releases it on every later failure
  DevBuf& o = s->bufs.back();
  if (kept && used == b.bytes && used >= ((size_t)32 << 20)) { // pinning and filling a second copy of a GB-sized array costs more than it saves
    o.src = v.data();
    CU(cudaMemcpy(o.d, o.src, o.bytes, cudaMemcpyHostToDevice));
  } else {
    CU(cudaMallocHost(&o.h, o.bytes));
    if (used) memcpy(o.h, v.data(), used);
    if (used < o.bytes) memset((char*)o.h + used, 0, o.bytes - used);
    CU(cudaMemcpy(o.d, o.h, o.bytes, cudaMemcpyHostToDevice));
  }
  *dptr = reinterpret_cast<const T*>(o.d);
  return SPT_OK;
}

// Grid of a persistent traversal kernel: `per_sm` resident blocks on every SM, fewer when there is little work.
int persistent_grid(const SptContext* c, long long n_rays, int per_sm) {
  const long long want = (n_rays + SPT_FETCH_CHUNK - 1) / SPT_FETCH_CHUNK; // one warp per chunk is enough
  const long long blocks = (want + (SPT_TRAV_BLOCK / 32) - 1) / (SPT_TRAV_BLOCK / 32);
  return (int)std::max<long long>(1, std::min<long long>(blocks, (long long)c->sm_count * per_sm));
}
int grid_for(const SptContext* c, long long n, int per_sm) {
  long long blocks = (n + SPT_TRAV_BLOCK - 1) / SPT_TRAV_BLOCK;
  return (int)std::max<long long>(1, std::min<long long>(blocks, (long long)c->sm_count * per_sm));
}


} // namespace

extern "C" {

const char* spt_last_error(void) { return g_err.c_str(); }
int32_t spt_version(void) { return 100; }

int32_t spt_context_create(int32_t device_id, SptContext** out) {
  if (!out) return fail(SPT_ERR_INVALID, "spt_context_create: out is null");
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) return fail(SPT_ERR_CUDA, std::string("no CUDA device (this library has no CPU fallback): ") + cudaGetErrorString(e));
  if (device_id < 0 || device_id >= n) return fail(SPT_ERR_INVALID, "spt_context_create: bad device id");
  CU(cudaSetDevice(device_id));
  SptContext* c = new (std::nothrow) SptContext();
  if (!c) return fail(SPT_ERR_INVALID, "out of host memory");
  c->device = device_id;
  cudaDeviceProp p;
  CU(cudaGetDeviceProperties(&p, device_id));
  c->sm_count = p.multiProcessorCount;
  *out = c;
  return SPT_OK;
}
int32_t spt_context_destroy(SptContext* ctx) { delete ctx; return SPT_OK; }

int32_t spt_scene_create(SptContext* ctx, const SptSceneDesc* desc, SptScene** out) {
  if (!ctx || !desc || !out) return fail(SPT_ERR_INVALID, "spt_scene_create: null argument");
  CU(cudaSetDevice(ctx->device));
  SptScene* s = new (std::nothrow) SptScene();
  if (!s) return fail(SPT_ERR_INVALID, "out of host memory");
  s->ctx = ctx;
  const auto t_create = std::chrono::steady_clock::now();
  const std::string err = spt::flatten_scene(*desc, s->hs);
  if (!err.empty()) { delete s; return fail(SPT_ERR_INVALID, err); }
  spt::HostScene& h = s->hs;
  DScene& d = s->ds;
  memset(&d, 0, sizeof(d));
  int st;
#define UP(vec, field) if ((st = upload(s, vec, &d.field)) != SPT_OK) { spt_scene_destroy(s); return st; }
#define UPH(vec, field) if ((st = upload(s, vec, &d.field, true, true)) != SPT_OK) { spt_scene_destroy(s); return st; }
  {
    auto al = [](size_t b) { return (std::max<size_t>(b, 16) + 255) & ~(size_t)255; };
    s->hot_bytes = al(h.nodes.size() * sizeof(spt::CwNode)) + al(h.tlas_inst.size() * sizeof(int32_t)) + al(h.inst.size() * sizeof(DInstance)) +
                   al(h.spheres.size() * sizeof(float4)) + al(h.tris.size() * sizeof(float4));
    if (cudaMalloc(&s->hot_arena, s->hot_bytes) != cudaSuccess) { cudaGetLastError(); s->hot_arena = nullptr; s->hot_bytes = 0; } // fall back to separate allocations
    std::vector<float4> nodes4(h.nodes.size() * 5);
    if (!h.nodes.empty()) memcpy(nodes4.data(), h.nodes.data(), h.nodes.size() * sizeof(spt::CwNode));
    if ((st = upload(s, nodes4, &d.nodes, false, true)) != SPT_OK) { spt_scene_destroy(s); return st; }
  }
  UPH(h.tlas_inst, tlas_inst); UPH(h.inst, inst); UPH(h.spheres, spheres); UPH(h.tris, tris); // nodes first: a window smaller than the arena keeps the tree
  UP(h.shade_tris, shade_tris); UP(h.inst_mat, inst_mat); UP(h.inst_shade, inst_shade);
  UP(h.meshes, meshes); UP(h.verts, verts); UP(h.indices, indices); UP(h.surfaces, surfaces); UP(h.materials, materials);
  UP(h.textures, textures); UP(h.mips, mips); UP(h.texels, texels); UP(h.mediums, mediums); UP(h.lights, lights);
  UP(h.light_prob, light_prob); UP(h.light_alias, light_alias); UP(h.inst_light, inst_light); UP(h.ss_cdf, ss_cdf);
  if (h.env_w > 0) { UP(h.env_tex, env_tex); UP(h.env_alias, env_alias); }
#undef UP
#undef UPH
  for (int c = 0; c < 3; c++) d.env_scale[c] = h.env_scale[c];
  d.env_w = h.env_w; d.env_h = h.env_h;
  d.n_lights = (int32_t)h.lights.size(); d.light_sampler = h.light_sampler; d.env_light_index = h.env_light_index;
  d.n_instances = (int32_t)h.inst.size(); d.tlas_empty = h.nodes.empty() || h.inst.empty();
  d.has_image_tex = h.has_image_tex ? 1 : 0;
  d.mat_mask = h.mat_mask; d.has_medium = h.has_medium ? 1 : 0;
  d.num_light_inv = 1.0f / (float)h.lights.size(); // UniformLightSampler::new, uniform.rs:19
  s->cameras = h.cameras;
  d.camera = s->cameras[0];
  // from here on a failed CUDA call must release what the scene already owns (uploads, streams, events)
#define CUS(call)                                                                                                      \
  do {                                                                                                                 \
    cudaError_t e_ = (call);                                                                                           \
    if (e_ != cudaSuccess) { spt_scene_destroy(s); return fail(SPT_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } \
  } while (0)
  CUS(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
  CUS(cudaEventCreate(&s->ev[0])); CUS(cudaEventCreate(&s->ev[1]));
  for (int i = 0; i < 4; i++) CUS(cudaEventCreate(&s->sev[i]));
  for (int i = 0; i < 3; i++) CUS(cudaStreamCreateWithFlags(&s->aux[i], cudaStreamNonBlocking));
  CUS(cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming)); CUS(cudaEventCreateWithFlags(&s->ev_shaded, cudaEventDisableTiming));
  CUS(cudaEventCreateWithFlags(&s->ev_con, cudaEventDisableTiming));
  for (int i = 0; i < 2; i++) CUS(cudaEventCreateWithFlags(&s->ev_join[i], cudaEventDisableTiming));
  CUS(cudaMalloc(&s->ctr, sizeof(DevCounters))); CUS(cudaMalloc(&s->batch_next, SPT_BATCH_CURSORS * sizeof(uint32_t)));
  CUS(cudaMallocHost(&s->h_ctr, sizeof(DevCounters)));
#undef CUS
  // Optional (SPT_L2_PERSIST=1): an L2 persisting access-policy window over the traversal arrays on the streams that run the
  // traversal kernels, so that the path pool streaming through the same L2 cannot evict the BVH.  Measured: see profiles/README.md.
  if (const char* e = getenv("SPT_L2_PERSIST")) if (atoi(e) > 0 && s->hot_arena) {
    int max_win = 0, max_persist = 0;
    cudaDeviceGetAttribute(&max_win, cudaDevAttrMaxAccessPolicyWindowSize, ctx->device);
    cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, ctx->device);
    const size_t win = std::min<size_t>(std::min<size_t>(s->hot_used, (size_t)max_win), (size_t)max_persist);
    if (win > 0 && cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, win) == cudaSuccess) {
      cudaStreamAttrValue av; memset(&av, 0, sizeof(av));
      av.accessPolicyWindow.base_ptr = s->hot_arena; av.accessPolicyWindow.num_bytes = win; av.accessPolicyWindow.hitRatio = 1.0f;
      av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting; av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
      bool ok = cudaStreamSetAttribute(s->stream, cudaStreamAttributeAccessPolicyWindow, &av) == cudaSuccess;
      for (int i = 0; i < 3; i++) ok = ok && cudaStreamSetAttribute(s->aux[i], cudaStreamAttributeAccessPolicyWindow, &av) == cudaSuccess;
      s->l2_window = ok;
    }
    cudaGetLastError();
  }
  s->hs.seconds_create = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_create).count();
  *out = s;
  return SPT_OK;
}

int32_t spt_scene_upload(SptScene* s, int64_t* bytes_copied) {
  if (!s) return fail(SPT_ERR_INVALID, "spt_scene_upload: null scene");
  CU(cudaSetDevice(s->ctx->device));
  int64_t total = 0;
  for (const DevBuf& b : s->bufs) { CU(cudaMemcpyAsync(b.d, b.h ? b.h : b.src, b.bytes, cudaMemcpyHostToDevice, s->stream)); total += (int64_t)b.bytes; }
  CU(cudaStreamSynchronize(s->stream));
  if (bytes_copied) *bytes_copied = total;
  return SPT_OK;
}

int32_t spt_scene_info(const SptScene* s, SptSceneInfo* out) {
  if (!s || !out) return fail(SPT_ERR_INVALID, "spt_scene_info: null argument");
  out->tlas_nodes = s->hs.tlas_nodes; out->blas_nodes = s->hs.blas_nodes;
  out->triangles = (int64_t)(s->hs.tris.size() / 3); out->instances = (int64_t)s->hs.inst.size();
  out->bytes_nodes = (int64_t)(s->hs.nodes.size() * sizeof(spt::CwNode)); out->bytes_tris = (int64_t)(s->hs.tris.size() * sizeof(float4));
  int64_t t = 0; for (const DevBuf& b : s->bufs) t += (int64_t)b.bytes;
  out->bytes_total = t;
  out->seconds_bvh = s->hs.seconds_bvh; out->seconds_create = s->hs.seconds_create; out->blas_on_device = s->hs.blas_on_device;
  return SPT_OK;
}

int32_t spt_scene_destroy(SptScene* s) {
  if (!s) return SPT_OK;
  cudaSetDevice(s->ctx->device);
  for (DevBuf& b : s->bufs) { if (!b.in_arena) cudaFree(b.d); if (b.h) cudaFreeHost(b.h); }
  cudaFree(s->hot_arena);
  cudaFree(s->pool_mem); cudaFree(s->film); cudaFree(s->ctr); cudaFree(s->batch_next); cudaFree(s->tiles);
  cudaFree(s->out_mean); cudaFree(s->out_rgb8); cudaFree(s->stage_rays); cudaFree(s->stage_out); cudaFree(s->gather_buf); cudaFree(s->all_tiles);
  if (s->h_ctr) cudaFreeHost(s->h_ctr);
  if (s->ev[0]) cudaEventDestroy(s->ev[0]);
  if (s->ev[1]) cudaEventDestroy(s->ev[1]);
  for (int i = 0; i < 4; i++) if (s->sev[i]) cudaEventDestroy(s->sev[i]);
  for (int i = 0; i < 3; i++) if (s->aux[i]) cudaStreamDestroy(s->aux[i]);
  if (s->ev_fork) cudaEventDestroy(s->ev_fork);
  if (s->ev_shaded) cudaEventDestroy(s->ev_shaded);
  if (s->ev_con) cudaEventDestroy(s->ev_con);
  for (int i = 0; i < 2; i++) if (s->ev_join[i]) cudaEventDestroy(s->ev_join[i]);
  if (s->stream) cudaStreamDestroy(s->stream);
  delete s;
  return SPT_OK;
}

// ---- seam B2 ---------------------------------------------------------------------------------------------------
static int32_t batch_device(SptScene* s, const SptRay* d_rays, int64_t n, void* d_out, bool any, uint32_t flags, SptStats* stats, void* stream_) {
  if (!s || (n > 0 && (!d_rays || !d_out)) || n < 0) return fail(SPT_ERR_INVALID, "ray batch: bad argument");
  CU(cudaSetDevice(s->ctx->device));
  // stream == NULL: the scene's own (non-blocking) stream, and the call returns only after the batch has finished — a caller
  // without a stream has nothing to order later work against.  With a caller stream the call is asynchronous.
  cudaStream_t st = stream_ ? (cudaStream_t)stream_ : s->stream;
  const bool count = (flags & SPT_RENDER_COUNT_TRAVERSAL) != 0;
  if (count) CU(cudaMemsetAsync(s->ctr, 0, sizeof(DevCounters), st));
  uint64_t launches = 0;
  const float4* r4 = reinterpret_cast<const float4*>(d_rays);
  const int64_t CH = 1ll << 30; // the fetch cursor of one launch is 32-bit
  for (int64_t off = 0; off < n; off += CH) {
    const uint32_t m = (uint32_t)std::min<int64_t>(CH, n - off);
    if (!s->occ_bat) { CU(spt::occupancy_traversal(&s->occ_ext, &s->occ_con, &s->occ_bat)); s->occ_bat = std::max(1, s->occ_bat); }
    const int grid = persistent_grid(s->ctx, m, s->occ_bat);
    // every launch takes the next cursor of the ring, zeroed on its own stream: batches issued on different caller streams may
    // overlap on the device without sharing a fetch cursor (up to SPT_BATCH_CURSORS launches in flight)
    uint32_t* cursor = s->batch_next + (s->batch_seq++ % SPT_BATCH_CURSORS);
    CU(cudaMemsetAsync(cursor, 0, sizeof(uint32_t), st));
    HitIO* ho = any ? nullptr : (HitIO*)d_out + off;
    uint8_t* oo = any ? (uint8_t*)d_out + off : nullptr;
    spt::launch_batch(s->ds, r4 + 2 * off, m, ho, oo, cursor, s->ctr, any, count, grid, st);
    CU(cudaGetLastError());
    launches++;
  }
  if (stats) {
    memset(stats, 0, sizeof(*stats));
    if (any) stats->shadow_rays = (uint64_t)n; else stats->ext_rays = (uint64_t)n;
    stats->kernel_launches = launches;
    if (count) {
      CU(cudaMemcpyAsync(s->h_ctr, s->ctr, sizeof(DevCounters), cudaMemcpyDeviceToHost, st));
      CU(cudaStreamSynchronize(st));
      stats->nodes_visited = s->h_ctr->nodes; stats->tris_tested = s->h_ctr->tris; stats->insts_visited = s->h_ctr->insts;
    }
  }
  if (!stream_) CU(cudaStreamSynchronize(st));
  return SPT_OK;
}
int32_t spt_trace_batch_device(SptScene* s, const SptRay* d_rays, int64_t n, SptHit* d_hits, uint32_t flags, SptStats* stats, void* stream) {
  return batch_device(s, d_rays, n, d_hits, false, flags, stats, stream);
}
int32_t spt_occluded_batch_device(SptScene* s, const SptRay* d_rays, int64_t n, uint8_t* d_out, uint32_t flags, SptStats* stats, void* stream) {
  return batch_device(s, d_rays, n, d_out, true, flags, stats, stream);
}
static int32_t batch_host(SptScene* s, const SptRay* rays, int64_t n, void* out, size_t out_elem, bool any) {
  if (!s || (n > 0 && (!rays || !out)) || n < 0) return fail(SPT_ERR_INVALID, "ray batch: bad argument");
  if (n == 0) return SPT_OK;
  CU(cudaSetDevice(s->ctx->device));
  // device staging kept by the scene and grown on demand: a per-frame B2 caller does not pay a cudaMalloc/cudaFree pair per call
  const size_t need_r = (size_t)n * sizeof(SptRay), need_o = (size_t)n * out_elem;
  if (s->stage_rays_bytes < need_r) { cudaFree(s->stage_rays); s->stage_rays = nullptr; s->stage_rays_bytes = 0; CU(cudaMalloc(&s->stage_rays, need_r)); s->stage_rays_bytes = need_r; }
  if (s->stage_out_bytes < need_o) { cudaFree(s->stage_out); s->stage_out = nullptr; s->stage_out_bytes = 0; CU(cudaMalloc(&s->stage_out, need_o)); s->stage_out_bytes = need_o; }
  SptRay* d_r = (SptRay*)s->stage_rays; void* d_o = s->stage_out;
  int32_t st = SPT_OK;
  if (cudaMemcpyAsync(d_r, rays, (size_t)n * sizeof(SptRay), cudaMemcpyHostToDevice, s->stream) != cudaSuccess) st = fail(SPT_ERR_CUDA, "H2D failed");
  if (st == SPT_OK) st = batch_device(s, d_r, n, d_o, any, 0, nullptr, s->stream);
  if (st == SPT_OK && cudaMemcpyAsync(out, d_o, (size_t)n * out_elem, cudaMemcpyDeviceToHost, s->stream) != cudaSuccess) st = fail(SPT_ERR_CUDA, "D2H failed");
  if (st == SPT_OK) { cudaError_t e = cudaStreamSynchronize(s->stream); if (e != cudaSuccess) st = fail(SPT_ERR_CUDA, std::string("ray batch: ") + cudaGetErrorString(e)); }
  return st;
}
int32_t spt_trace_batch(SptScene* s, const SptRay* rays, int64_t n, SptHit* out) { return batch_host(s, rays, n, out, sizeof(SptHit), false); }
int32_t spt_occluded_batch(SptScene* s, const SptRay* rays, int64_t n, uint8_t* out) { return batch_host(s, rays, n, out, 1, true); }

// No counterpart in the reference (see spt.h).  Allocates nothing: the scratch is a
// file-scope device array in trav_kernels.cu.
int32_t spt_trace_batch_noalloc(SptScene* s, const SptRay* rays, int64_t n, SptHit* out) {
  if (!s || n < 0 || (n > 0 && (!rays || !out))) return fail(SPT_ERR_INVALID, "spt_trace_batch_noalloc: bad argument");
  if (n > SPT_NOALLOC_MAX_RAYS)
    return fail(SPT_ERR_INVALID, "spt_trace_batch_noalloc: n = " + std::to_string(n) + " exceeds SPT_NOALLOC_MAX_RAYS = " + std::to_string(SPT_NOALLOC_MAX_RAYS) +
                                 " (the fixed scratch array); split the batch or use spt_trace_batch");
  if (n == 0) return SPT_OK;
  CU(cudaSetDevice(s->ctx->device));
  if (!s->occ_bat) { CU(spt::occupancy_traversal(&s->occ_ext, &s->occ_con, &s->occ_bat)); s->occ_bat = std::max(1, s->occ_bat); }
  const cudaError_t e = spt::trace_batch_noalloc(s->ds, rays, (uint32_t)n, out, persistent_grid(s->ctx, n, s->occ_bat), s->stream);
  if (e != cudaSuccess) return fail(SPT_ERR_CUDA, std::string("spt_trace_batch_noalloc: ") + cudaGetErrorString(e));
  return SPT_OK;
}

} // extern "C"

extern "C" int32_t spt_bxdf_batch(SptScene* s, const SptBxdfQuery* queries, int64_t n, SptBxdfResult* out) {
  if (!s || n < 0 || (n > 0 && (!queries || !out))) return fail(SPT_ERR_INVALID, "spt_bxdf_batch: bad argument");
  if (n == 0) return SPT_OK;
  for (int64_t i = 0; i < n; i++) if (queries[i].material < 0 || queries[i].material >= (int)s->hs.materials.size()) return fail(SPT_ERR_INVALID, "spt_bxdf_batch: bad material index");
  CU(cudaSetDevice(s->ctx->device));
  SptBxdfQuery* dq = nullptr; SptBxdfResult* dr = nullptr;
  CU(cudaMalloc(&dq, (size_t)n * sizeof(SptBxdfQuery)));
  if (cudaMalloc(&dr, (size_t)n * sizeof(SptBxdfResult)) != cudaSuccess) { cudaFree(dq); return fail(SPT_ERR_CUDA, "cudaMalloc failed"); }
  cudaError_t e = cudaMemcpyAsync(dq, queries, (size_t)n * sizeof(SptBxdfQuery), cudaMemcpyHostToDevice, s->stream);
  if (e == cudaSuccess) { spt::launch_bxdf(s->ds, dq, n, dr, grid_for(s->ctx, n, 8), s->stream); e = cudaGetLastError(); }
  if (e == cudaSuccess) e = cudaMemcpyAsync(out, dr, (size_t)n * sizeof(SptBxdfResult), cudaMemcpyDeviceToHost, s->stream);
  const cudaError_t e2 = cudaStreamSynchronize(s->stream); // always drain the stream before the buffers are freed
  if (e == cudaSuccess) e = e2;
  cudaFree(dq); cudaFree(dr);
  if (e != cudaSuccess) return fail(SPT_ERR_CUDA, std::string("spt_bxdf_batch: ") + cudaGetErrorString(e));
  return SPT_OK;
}

// ---- seam B1: RendererT::render ------------------------------------------------------------------------------------------
namespace {

uint32_t tile_hash(uint32_t t) { t ^= t >> 16; t *= 0x7feb352du; t ^= t >> 15; t *= 0x846ca68bu; t ^= t >> 16; return t; }

std::vector<uint32_t> owned_tiles(int width, int height, int tw, int th, int rank, int world) {
  const int tiles_x = (width + tw - 1) / tw, tiles_y = (height + th - 1) / th, n_all = tiles_x * tiles_y;
  std::vector<std::pair<uint32_t, uint32_t>> order(n_all);
  for (int t = 0; t < n_all; t++) order[t] = {tile_hash((uint32_t)t), (uint32_t)t};
  if (world > 1) std::sort(order.begin(), order.end());
  std::vector<uint32_t> mine;
  for (int i = rank; i < n_all; i += world) mine.push_back(order[i].second);
  std::sort(mine.begin(), mine.end());
  return mine;
}

// Film::filter_to_image on the device: the film of the last render -> mean image (+ 8-bit encode)
void resolve_film(SptScene* s, float radius, float* d_mean, uint8_t* d_rgb8, cudaStream_t st) {
  const RenderParams& P = s->last_P;
  const size_t n_px = (size_t)P.width * P.height;
  if (radius == 0.5f) // own-pixel mean (Q14: radius_int = 0 and every sample of the pixel lies in the box)
    spt::launch_resolve(s->film, (int)n_px, 1.0f / (float)s->film_spp, d_mean, d_rgb8, grid_for(s->ctx, (long long)n_px * 3, 8), st);
  else
    spt::launch_resolve_box(P, s->film_k0, s->film_spp, radius, (int)std::ceil(radius - 0.5f), d_mean, d_rgb8, grid_for(s->ctx, (long long)n_px, 16), st);
}

// d_mean == nullptr: leave the film unresolved (spt_render_gather resolves on the root after the exchange)
int render_impl(SptScene* s, const SptRenderDesc* r, float* d_mean, uint8_t* d_rgb8, SptStats* stats) {
  if (!s || !r) return fail(SPT_ERR_INVALID, "spt_render: null argument");
  if (r->width <= 0 || r->height <= 0 || r->max_depth < 0 || r->max_depth > 255) return fail(SPT_ERR_INVALID, "spt_render: bad width/height/max_depth");
  if (r->camera_index < 0 || r->camera_index >= (int)s->cameras.size()) return fail(SPT_ERR_INVALID, "spt_render: bad camera index");
  if (r->sampler_kind < 0 || r->sampler_kind > 2) return fail(SPT_ERR_INVALID, "spt_render: unknown sampler");
  if (!(r->filter_radius > 0.0f) || r->filter_radius > 3.5f) return fail(SPT_ERR_INVALID, "spt_render: box filter radius must be in (0, 3.5]");
  if (s->hs.lights.empty()) return fail(SPT_ERR_UNSUPPORTED, "spt_render: the scene has no light (the reference's light sampler would panic)");
  const int world = std::max(1, r->world_size), rank = r->rank;
  if (rank < 0 || rank >= world) return fail(SPT_ERR_INVALID, "spt_render: bad rank");
  const int spp = r->sampler_kind == SPT_SAMPLER_JITTERED ? std::max(1, r->div_x) * std::max(1, r->div_y) : std::max(1, r->spp);
  const int tw = r->tile_w > 0 ? r->tile_w : SPT_DEFAULT_TILE_W, th = r->tile_h > 0 ? r->tile_h : SPT_DEFAULT_TILE_H;
  if (tw % 8 || th % 4) return fail(SPT_ERR_INVALID, "spt_render: tile_w must be a multiple of 8 and tile_h of 4");
  // sample range of this call (progressive passes): [spp_first, spp_first + spp_pass) of the frame's spp
  const int spp_first = r->spp_count > 0 ? r->spp_first : 0, spp_pass = r->spp_count > 0 ? r->spp_count : spp;
  if (spp_first < 0 || spp_first + spp_pass > spp) return fail(SPT_ERR_INVALID, "spt_render: spp_first + spp_count exceeds the frame's spp");
  const bool accumulate = (r->flags & SPT_RENDER_ACCUMULATE) != 0 && s->film && s->film_w == r->width && s->film_h == r->height && s->film_spp > 0;
  if ((r->flags & SPT_RENDER_ACCUMULATE) && !accumulate && spp_first > 0)
    return fail(SPT_ERR_INVALID, "spt_render: SPT_RENDER_ACCUMULATE with spp_first > 0 needs the film of a previous pass of the same size (or spt_film_load)");
  CU(cudaSetDevice(s->ctx->device));
  cudaStream_t st = s->stream;
  const size_t n_px = (size_t)r->width * r->height;
  // ---- buffers
  // default pool: sized to this rank's camera samples, between 2^22 and 2^25 slots (fewer for renders smaller than 2^22 samples).
  // Every iteration of the wavefront loop pays ~110 us of kernel tails and launch gaps whatever the pool holds, so long renders
  // want a big pool (C2..C5 at full length: 2^23 +5 %, 2^24 +7..10 % over 2^22; profiles/README.md), while a render of a few
  // pool-fills spends its time filling and draining the pool and wants a small one: the measured optimum is a pool that
  // finishes the render in ~20 iterations.  How many samples an iteration starts depends on how long the scene's paths
  // live, which the first iterations tell: the render starts on samples/6 slots (samples/16 in round 1, right for C1 then, whose paths mostly end at
  // their first hit) and, after three iterations, widens the ACTIVE part of the pool up to the allocated samples/4 (aiming at ~10 iterations).
  // 2^25 slots = 6.0 GB (the cap was 2^24 in round 1, when 2^25 gained 0.6 %; with the faster kernels of round 2 it gains 1.0 % on C2 and
  // 1.6 % on C3, 2^26 nothing more; 6 GB of a 180 GB device).  SPT_POOL_CAP_LOG2 overrides it.
  const unsigned long long all_samples = (unsigned long long)n_px * (unsigned long long)spp_pass;
  const unsigned long long pool_lo = std::min<unsigned long long>(1ull << 22, std::max<unsigned long long>(1ull << 16, all_samples));
  const unsigned long long rank_samples = all_samples / (unsigned long long)world;
  auto round_pool = [](unsigned long long n) { return (int)std::max<unsigned long long>(SPT_TRAV_BLOCK, (n + SPT_TRAV_BLOCK - 1) / SPT_TRAV_BLOCK * SPT_TRAV_BLOCK); };
  // SPT_POOL_CAP_DIV / SPT_POOL_ITERS: the two constants of the rule (allocated = samples / cap_div; widening aims at `iters` iterations), for sweeps
  unsigned long long cap_div = 4ull; double want_iters = 10.0; // re-tuned in round 2 (8 / 20 before): the kernels got 25 % faster, the per-iteration tails did not
  if (const char* e = getenv("SPT_POOL_CAP_DIV")) cap_div = std::max(1, atoi(e));
  if (const char* e = getenv("SPT_POOL_ITERS")) want_iters = std::max(1.0, atof(e));
  unsigned long long start_div = 6ull; // 16 in round 1; 6 measured best on renders of 60-130 M samples (C3 +2.8 %, C4 +2.2 %, C5 +1.7 %, C1 +0.8 %, C2 0 %)
  if (const char* e = getenv("SPT_POOL_START_DIV")) start_div = std::max(1, atoi(e));
  unsigned long long pool_cap = 1ull << 25;
  if (const char* e = getenv("SPT_POOL_CAP_LOG2")) pool_cap = 1ull << std::min(28, std::max(16, atoi(e)));
  int pool_n = r->pool_size > 0 ? r->pool_size : (int)std::min<unsigned long long>(pool_cap, std::max<unsigned long long>(pool_lo, rank_samples / cap_div)); // allocated
  int pool_start = r->pool_size > 0 ? r->pool_size : (int)std::min<unsigned long long>(pool_cap, std::max<unsigned long long>(pool_lo, rank_samples / start_div)); // active at first
  pool_n = round_pool(pool_n); pool_start = std::min(pool_n, round_pool(pool_start));
  const size_t per_slot = sizeof(PathRec) + sizeof(float4) + 3 * sizeof(float4) + sizeof(int32_t) + sizeof(uint32_t) + (SPT_NQ + 2) * sizeof(uint32_t);
  if (s->pool_n != pool_n) {
    cudaFree(s->pool_mem); s->pool_mem = nullptr; s->pool_n = 0;
    // a pool the CALLER asked for must fit; the default one shrinks (down to the 2^22 rule) if the device is short of memory
    while (r->pool_size <= 0 && cudaMalloc(&s->pool_mem, per_slot * (size_t)pool_n) != cudaSuccess) {
      cudaGetLastError(); s->pool_mem = nullptr;
      if ((unsigned long long)pool_n / 2 < pool_lo) break;
      pool_n = round_pool(pool_n / 2); pool_start = std::min(pool_start, pool_n);
    }
    if (!s->pool_mem) CU(cudaMalloc(&s->pool_mem, per_slot * (size_t)pool_n));
    s->pool_n = pool_n; s->pool_bytes = per_slot * (size_t)pool_n;
  }
  if (s->film_px < n_px) { cudaFree(s->film); s->film = nullptr; s->film_spp = 0; CU(cudaMalloc(&s->film, n_px * 3 * sizeof(long long))); s->film_px = n_px; }
  RenderParams P; memset(&P, 0, sizeof(P));
  {
    char* p = (char*)s->pool_mem; const size_t N = (size_t)pool_n;
    P.pool.rec = (PathRec*)p; p += N * sizeof(PathRec);
    P.pool.hit = (float4*)p; p += N * sizeof(float4);
    P.sq.o = (float4*)p; p += N * sizeof(float4);
    P.sq.d = (float4*)p; p += N * sizeof(float4);
    P.sq.c = (float4*)p; p += N * sizeof(float4);
    P.pool.hit_inst = (int32_t*)p; p += N * sizeof(int32_t);
    P.pool.flags = (uint32_t*)p; p += N * sizeof(uint32_t);
    P.queue = (uint32_t*)p; p += N * SPT_NQ * sizeof(uint32_t);
    P.ext_queue[0] = (uint32_t*)p; p += N * sizeof(uint32_t);
    P.ext_queue[1] = (uint32_t*)p; p += N * sizeof(uint32_t);
    P.pool.n = pool_start; // active slots; the arrays hold pool_n
  }
  P.ctr = s->ctr; P.film = s->film;
  // ---- tile ownership: tiles sorted by a hash, dealt round-robin -> interleaved, balanced to +-1 tile
  const int tiles_x = (r->width + tw - 1) / tw;
  const int tile_key[6] = {r->width, r->height, tw, th, rank, world};
  if (!s->tiles || memcmp(tile_key, s->tile_key, sizeof(tile_key)) != 0) { // the list only depends on these six numbers: keep it across renders
    const std::vector<uint32_t> mine = owned_tiles(r->width, r->height, tw, th, rank, world);
    if (s->tiles_cap < mine.size() || !s->tiles) { cudaFree(s->tiles); s->tiles = nullptr; CU(cudaMalloc(&s->tiles, std::max<size_t>(1, mine.size()) * sizeof(uint32_t))); s->tiles_cap = std::max<size_t>(1, mine.size()); }
    if (!mine.empty()) CU(cudaMemcpy(s->tiles, mine.data(), mine.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    s->n_tiles = mine.size();
    memcpy(s->tile_key, tile_key, sizeof(tile_key));
  }
  P.tiles = s->tiles; P.n_tiles = (int)s->n_tiles; P.tiles_x = tiles_x; P.tile_w = tw; P.tile_h = th;
  P.width = r->width; P.height = r->height; P.max_depth = r->max_depth;
  P.sampler = r->sampler_kind; P.spp = spp_pass; P.spp_first = spp_first; P.spp_frame = spp; P.div_x = std::max(1, r->div_x); P.div_y = std::max(1, r->div_y);
  P.seed = r->seed;
  P.sort_materials = (r->flags & SPT_RENDER_NO_SORT) ? 0 : 1;
  { const char* e = getenv("SPT_EXT_QUEUE"); P.use_ext_queue = (e && atoi(e) == 0) ? 0 : 1; }
  P.total_samples = (unsigned long long)s->n_tiles * (unsigned long long)(tw * th) * (unsigned long long)spp_pass;
  DScene ds = s->ds;
  ds.camera = s->cameras[r->camera_index];
  const bool count = (r->flags & SPT_RENDER_COUNT_TRAVERSAL) != 0;
  // ---- run
  if (!accumulate) { CU(cudaMemsetAsync(s->film, 0, n_px * 3 * sizeof(long long), st)); s->film_spp = 0; s->film_k0 = spp_first; s->film_contiguous = true; }
  else if (spp_first != s->film_k0 + s->film_spp) s->film_contiguous = false;
  // a radius other than 0.5 reconstructs from the sub-pixel offsets of the accumulated samples, recomputed from (pixel, k): the film
  // must hold one consecutive range of sample indices (rendered with this call's seed and sampler)
  if (r->filter_radius != 0.5f && !s->film_contiguous)
    return fail(SPT_ERR_UNSUPPORTED, "spt_render: a box filter radius other than 0.5 needs progressive passes over consecutive sample ranges");
  s->film_w = r->width; s->film_h = r->height; s->film_spp += spp_pass;
  CU(cudaMemsetAsync(s->ctr, 0, sizeof(DevCounters), st));
  CU(cudaMemsetAsync(P.pool.flags, 0, (size_t)pool_n * sizeof(uint32_t), st));
  CU(cudaMemsetAsync(P.pool.hit_inst, 0xff, (size_t)pool_n * sizeof(int32_t), st));
  CU(cudaEventRecord(s->ev[0], st));
  if (!s->occ_ext) { // resident blocks per SM of the persistent kernels, from the occupancy API (register / shared-memory limits)
    CU(spt::occupancy_traversal(&s->occ_ext, &s->occ_con, &s->occ_bat));
    s->occ_bat = std::max(1, s->occ_bat);
    s->occ_ext = std::max(1, s->occ_ext); s->occ_con = std::max(1, s->occ_con);
  }
  int grid_ext = 0, grid_con = 0, grid_shade = 0, grid_shade_lean = 0, grid_logic = 0, grid_gen = 0;
  auto set_active_pool = [&](int n) { // between iterations only: the queues are empty then, and slots never used are free slots
    P.pool.n = n;
    grid_ext = persistent_grid(s->ctx, n, s->occ_ext); grid_con = persistent_grid(s->ctx, n, s->occ_con);
    grid_shade = grid_for(s->ctx, n, 4); // = the resident blocks per SM (ShadeCfg::min_blocks): ONE wave of grid-stride blocks.  16 per SM cost 0.7 % (C2) .. 3.5 % (C3), 5 or 6 (a ragged second wave) 5-20 % of the shade stage
    grid_shade_lean = grid_for(s->ctx, n, SPT_SHADE_LEAN_MINB);
    grid_logic = (n + SPT_LOGIC_BLOCK - 1) / SPT_LOGIC_BLOCK;
    grid_gen = std::min(grid_logic, s->ctx->sm_count * 8);
  };
  set_active_pool(pool_start);
  uint64_t iters = 0, launches = 0;
  const bool timed = (r->flags & SPT_RENDER_TIME_STAGES) != 0;
  double stage[6] = {0, 0, 0, 0, 0, 0}; // generate (inside shade), extend, shade (logic..generate), connect, resolve, other
  // SPT_TRACE_ITERS=file (with SPT_RENDER_TIME_STAGES): one line per iteration — iteration, active pool, live paths, sample cursor,
  // shadow rays, ms of logic..generate / connect / extend — the fill / steady / drain profile of a render (tools/share_profile.py)
  FILE* trace = nullptr;
  if (timed) if (const char* e = getenv("SPT_TRACE_ITERS")) trace = fopen(e, "a");
  const uint32_t mats = P.sort_materials ? ds.mat_mask : 0u;
  // scenes whose textures are all plain scalars run the lean compilation of the shade kernels (SPT_SHADE_LEAN=0 forces the general one)
  bool lean = s->hs.lean;
  if (const char* e = getenv("SPT_SHADE_LEAN")) lean = lean && atoi(e) != 0;
  const bool special = ds.has_medium || ds.env_tex != nullptr;
  // Within one iteration the shade kernels of different material classes are independent of each other, and connect is
  // independent of generate + extend: they are issued on separate streams so that small kernels overlap and the tail of
  // one persistent kernel is filled by the blocks of the next (serial = one stream, used for per-stage timing).
  const bool serial = timed || (r->flags & SPT_RENDER_SERIAL) != 0;
  bool con_pending = false;
  const bool aov_normal = (r->flags & SPT_RENDER_DEBUG_NORMAL) != 0;
  auto iteration = [&]() {
    P.parity = (int)(iters & 1u); // which of the two live-slot lists this iteration writes (wavefront_types.h)
    if (aov_normal) { // pt.rs:113-118: every path ends at its first hit; no shading, no shadow rays
      spt::launch_aov_normal(ds, P, grid_gen, st);
      spt::launch_generate(ds, P, grid_gen, st);
      spt::launch_extend(ds, P, count, grid_ext, st);
      launches += 3; iters++;
      return;
    }
    if (con_pending) { cudaStreamWaitEvent(st, s->ev_con, 0); con_pending = false; } // logic resets the shadow-queue counters
    spt::launch_logic(ds, P, grid_logic, st); launches++;
    if (special) { spt::launch_special(ds, P, grid_shade, st); launches++; }
    if (serial) {
      for (int m = 0; m < 8; m++) if (mats & (1u << m)) { spt::launch_shade(m, ds, P, count, lean, lean ? grid_shade_lean : grid_shade, st); launches++; }
      if (!P.sort_materials) { spt::launch_shade(-1, ds, P, count, false, grid_shade, st); launches++; }
      spt::launch_generate(ds, P, grid_gen, st); launches++;
      if (timed) cudaEventRecord(s->sev[1], st);
      spt::launch_connect(ds, P, count, grid_con, st);
      if (timed) cudaEventRecord(s->sev[2], st);
      spt::launch_extend(ds, P, count, grid_ext, st);
    } else {
      cudaEventRecord(s->ev_fork, st);
      int k = 0; bool used[2] = {false, false};
      for (int m = 0; m < 8; m++) if (mats & (1u << m)) {
        cudaStream_t sm = st;
        if (k > 0) { const int j = (k - 1) & 1; sm = s->aux[j]; if (!used[j]) { cudaStreamWaitEvent(sm, s->ev_fork, 0); used[j] = true; } }
        spt::launch_shade(m, ds, P, count, lean, lean ? grid_shade_lean : grid_shade, sm); launches++; k++;
      }
      if (!P.sort_materials) { spt::launch_shade(-1, ds, P, count, false, grid_shade, st); launches++; }
      for (int j = 0; j < 2; j++) if (used[j]) { cudaEventRecord(s->ev_join[j], s->aux[j]); cudaStreamWaitEvent(st, s->ev_join[j], 0); }
      cudaEventRecord(s->ev_shaded, st);
      cudaStreamWaitEvent(s->aux[2], s->ev_shaded, 0);
      spt::launch_connect(ds, P, count, grid_con, s->aux[2]);
      cudaEventRecord(s->ev_con, s->aux[2]); con_pending = true;
      spt::launch_generate(ds, P, grid_gen, st); launches++;
      spt::launch_extend(ds, P, count, grid_ext, st);
    }
    launches += 2; iters++;
  };
  if ((r->max_depth > 0 || aov_normal) && P.total_samples > 0) {
    // The host only needs the device counters to know when to stop.  While camera samples remain, it looks every
    // `stride` iterations (the cursor tells how far the render is); once the cursor is exhausted the pool drains
    // within max_depth iterations and the host looks after every one.
    bool drain_seen = false;
    unsigned long long cursor_1 = 0; // cursor after the first iteration (= the pool filled once)
    bool widen = !timed && pool_n > pool_start;
    for (;;) {
      int stride = 1;
      if (widen && iters == 1) { cursor_1 = s->h_ctr->cursor; stride = 2; }
      else if (!timed && iters > 0 && s->h_ctr->cursor < P.total_samples) {
        const double done = (double)s->h_ctr->cursor / (double)P.total_samples;
        const double per_iter = done / (double)iters;                       // fraction of the samples started per iteration
        const double left = per_iter > 0 ? (1.0 - done) / per_iter : 1.0;   // iterations until the cursor runs out
        stride = (int)std::max(1.0, std::min(64.0, left * 0.5));
      } else if (!timed && iters > 0 && !drain_seen) {
        // first look after the last camera sample started: a path lives at most max_depth iterations, so issue most of
        // the drain back to back (iterations over an empty pool cost a few tens of microseconds) instead of idling the
        // GPU on a host round trip after each one
        drain_seen = true;
        stride = std::max(1, r->max_depth - 2);
      }
      if (widen && iters == 3) { // samples started per iteration and slot, now that the pool has turned over twice
        widen = false;
        const double per_slot_iter = (double)(s->h_ctr->cursor - cursor_1) / (2.0 * (double)P.pool.n);
        if (per_slot_iter > 0 && s->h_ctr->cursor < P.total_samples) {
          const double want = (double)P.total_samples / (want_iters * per_slot_iter);
          set_active_pool(std::min(pool_n, std::max(P.pool.n, round_pool((unsigned long long)want))));
        }
      }
      for (int k = 0; k < stride; k++) {
        if (timed) cudaEventRecord(s->sev[0], st);
        iteration();
        if (timed) cudaEventRecord(s->sev[3], st);
      }
      if (con_pending) { cudaStreamWaitEvent(st, s->ev_con, 0); con_pending = false; }
      CU(cudaMemcpyAsync(s->h_ctr, s->ctr, sizeof(DevCounters), cudaMemcpyDeviceToHost, st));
      CU(cudaStreamSynchronize(st));
      CU(cudaGetLastError());
      if (timed) {
        float a = 0, b = 0, c = 0;
        cudaEventElapsedTime(&a, s->sev[0], s->sev[1]); cudaEventElapsedTime(&b, s->sev[1], s->sev[2]); cudaEventElapsedTime(&c, s->sev[2], s->sev[3]);
        stage[2] += a * 1e-3; stage[3] += b * 1e-3; stage[1] += c * 1e-3;
        if (trace) fprintf(trace, "%llu %d %u %llu %u %.4f %.4f %.4f\n", (unsigned long long)iters, P.pool.n, s->h_ctr->it.live, s->h_ctr->cursor, s->h_ctr->it.sq, a, b, c);
      }
      // done = nothing in flight AND every camera sample started: an iteration can end with no live path while samples remain
      // (all of its ordinals fell into the padding of edge tiles, or max_depth ended every path it started)
      if (s->h_ctr->it.live == 0 && s->h_ctr->cursor >= P.total_samples) break;
    }
  }
  if (trace) fclose(trace);
  s->last_P = P;
  if (d_mean) { resolve_film(s, r->filter_radius, d_mean, d_rgb8, st); launches++; }
  CU(cudaEventRecord(s->ev[1], st));
  CU(cudaMemcpyAsync(s->h_ctr, s->ctr, sizeof(DevCounters), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  CU(cudaGetLastError());
  if (stats) {
    memset(stats, 0, sizeof(*stats));
    float ms = 0.f; cudaEventElapsedTime(&ms, s->ev[0], s->ev[1]);
    const DevCounters& c = *s->h_ctr;
    stats->paths = c.paths; stats->ext_rays = c.ext_rays; stats->shadow_rays = c.shadow_rays; stats->probe_rays = c.probe_rays;
    stats->nodes_visited = c.nodes; stats->tris_tested = c.tris; stats->insts_visited = c.insts; stats->nonfinite = c.nonfinite; stats->shadow_culled = c.shadow_culled;
    for (int i = 0; i < 10; i++) stats->trav_profile[i] = c.trav_profile[i];
    stats->iterations = iters; stats->kernel_launches = launches; stats->seconds_total = ms * 1e-3;
    for (int i = 0; i < 6; i++) stats->seconds_stage[i] = stage[i];
  }
  return SPT_OK;
}

} // namespace

extern "C" int32_t spt_film_save(SptScene* s, int64_t* out_values, int64_t n_values) {
  if (!s || !out_values) return fail(SPT_ERR_INVALID, "spt_film_save: null argument");
  if (!s->film || s->film_spp <= 0) return fail(SPT_ERR_INVALID, "spt_film_save: no film (render first)");
  if (n_values != (int64_t)s->film_w * s->film_h * 3) return fail(SPT_ERR_INVALID, "spt_film_save: n_values must be width*height*3 of the last render");
  CU(cudaSetDevice(s->ctx->device));
  CU(cudaMemcpy(out_values, s->film, (size_t)n_values * sizeof(long long), cudaMemcpyDeviceToHost));
  return SPT_OK;
}

extern "C" int32_t spt_film_load(SptScene* s, const int64_t* values, int32_t width, int32_t height, int32_t spp_done) {
  if (!s || !values) return fail(SPT_ERR_INVALID, "spt_film_load: null argument");
  if (width <= 0 || height <= 0 || spp_done <= 0) return fail(SPT_ERR_INVALID, "spt_film_load: bad width/height/spp_done");
  CU(cudaSetDevice(s->ctx->device));
  const size_t n_px = (size_t)width * height;
  if (s->film_px < n_px) { cudaFree(s->film); s->film = nullptr; s->film_spp = 0; CU(cudaMalloc(&s->film, n_px * 3 * sizeof(long long))); s->film_px = n_px; }
  CU(cudaMemcpy(s->film, values, n_px * 3 * sizeof(long long), cudaMemcpyHostToDevice));
  s->film_w = width; s->film_h = height; s->film_spp = spp_done; s->film_k0 = 0; s->film_contiguous = true; // a checkpoint holds samples [0, spp_done)
  return SPT_OK;
}

extern "C" int32_t spt_render_device(SptScene* s, const SptRenderDesc* r, float* d_rgb_mean, SptStats* stats) {
  if (!d_rgb_mean) return fail(SPT_ERR_INVALID, "spt_render_device: null output");
  return render_impl(s, r, d_rgb_mean, nullptr, stats);
}

// ---- multi-GPU: the rank join and the framebuffer exchange behind the C ABI -------------------------------------------------------------
// NCCL is bound at run time (dlopen of libnccl.so.2: the copy a host such as PyTorch has already loaded, else the system one), so the
// library has no link-time dependency on it and single-GPU hosts never touch it.
namespace {
struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;
std::mutex g_nccl_mutex;
int nccl_load() {
  std::lock_guard<std::mutex> lock(g_nccl_mutex); // two hosts threads may make their first spt_comm_* call at once
  if (g_nccl.lib) return SPT_OK;
  // A process holds ONE libnccl.so.2 (the loader matches by soname): if the host also uses another NCCL consumer with its own copy
  // (PyTorch bundles a newer one than the system's), that consumer must be loaded FIRST, or it will bind to whatever was loaded
  // here and may miss symbols.  SPT_NCCL_LIB names the file to load instead of the default search.
  void* h = nullptr;
  if (const char* e = getenv("SPT_NCCL_LIB")) h = dlopen(e, RTLD_NOW | RTLD_LOCAL);
  for (const char* name : {"libnccl.so.2", "libnccl.so"}) { if (h) break; h = dlopen(name, RTLD_NOW | RTLD_LOCAL); }
  if (!h) return fail(SPT_ERR_UNSUPPORTED, std::string("spt_comm: cannot load libnccl.so.2 (") + dlerror() + ")");
  NcclApi a; a.lib = h;
  bool ok = true;
  auto sym = [&](const char* n) { void* p = dlsym(h, n); ok = ok && p; return p; };
  a.GetUniqueId = (decltype(a.GetUniqueId))sym("ncclGetUniqueId"); a.CommInitRank = (decltype(a.CommInitRank))sym("ncclCommInitRank");
  a.CommDestroy = (decltype(a.CommDestroy))sym("ncclCommDestroy"); a.Send = (decltype(a.Send))sym("ncclSend"); a.Recv = (decltype(a.Recv))sym("ncclRecv");
  a.GroupStart = (decltype(a.GroupStart))sym("ncclGroupStart"); a.GroupEnd = (decltype(a.GroupEnd))sym("ncclGroupEnd");
  a.GetErrorString = (decltype(a.GetErrorString))sym("ncclGetErrorString");
  if (!ok) return fail(SPT_ERR_UNSUPPORTED, "spt_comm: libnccl.so.2 lacks a required symbol");
  g_nccl = a;
  return SPT_OK;
}
#define NC(call)                                                                                                   \
  do {                                                                                                             \
    ncclResult_t r_ = (call);                                                                                      \
    if (r_ != ncclSuccess) return fail(SPT_ERR_CUDA, std::string(#call) + ": " + g_nccl.GetErrorString(r_));        \
  } while (0)
} // namespace

struct SptComm { SptContext* ctx = nullptr; ncclComm_t comm = nullptr; int rank = 0, world = 1; };

extern "C" int32_t spt_comm_unique_id(uint8_t* out_id) {
  if (!out_id) return fail(SPT_ERR_INVALID, "spt_comm_unique_id: null argument");
  static_assert(sizeof(ncclUniqueId) == SPT_COMM_ID_BYTES, "SPT_COMM_ID_BYTES must equal sizeof(ncclUniqueId)");
  const int st = nccl_load();
  if (st != SPT_OK) return st;
  ncclUniqueId id;
  NC(g_nccl.GetUniqueId(&id));
  memcpy(out_id, &id, sizeof(id));
  return SPT_OK;
}

extern "C" int32_t spt_comm_create(SptContext* ctx, const uint8_t* id, int32_t rank, int32_t world_size, SptComm** out) {
  if (!ctx || !id || !out || world_size < 1 || rank < 0 || rank >= world_size) return fail(SPT_ERR_INVALID, "spt_comm_create: bad argument");
  const int st = nccl_load();
  if (st != SPT_OK) return st;
  CU(cudaSetDevice(ctx->device));
  ncclUniqueId nid; memcpy(&nid, id, sizeof(nid));
  SptComm* c = new (std::nothrow) SptComm();
  if (!c) return fail(SPT_ERR_INVALID, "out of host memory");
  c->ctx = ctx; c->rank = rank; c->world = world_size;
  const ncclResult_t r = g_nccl.CommInitRank(&c->comm, world_size, nid, rank);
  if (r != ncclSuccess) { delete c; return fail(SPT_ERR_CUDA, std::string("ncclCommInitRank: ") + g_nccl.GetErrorString(r)); }
  *out = c;
  return SPT_OK;
}

extern "C" int32_t spt_comm_destroy(SptComm* c) {
  if (!c) return SPT_OK;
  if (c->comm && g_nccl.CommDestroy) { cudaSetDevice(c->ctx->device); g_nccl.CommDestroy(c->comm); }
  delete c;
  return SPT_OK;
}

// Render this rank's tiles, then move ONLY the tiles each rank owns — as the integer film, 3 x int64 per pixel — to `root` with
// grouped ncclSend / ncclRecv, unpack them into the root's film and reconstruct there.  W*H*24 bytes cross NVLink in total instead of
// a W*H*12-byte full-frame reduce PER RANK, and because the film is integer the gathered frame is bit-identical to the one-GPU frame
// for every filter radius (the neighbour sums a radius > 0.5 needs exist on the root after the unpack).
extern "C" int32_t spt_render_gather(SptScene* s, SptComm* c, const SptRenderDesc* desc, int32_t root, float* d_rgb_mean, SptStats* stats) {
  if (!s || !c || !desc) return fail(SPT_ERR_INVALID, "spt_render_gather: null argument");
  if (root < 0 || root >= c->world) return fail(SPT_ERR_INVALID, "spt_render_gather: bad root");
  if (c->rank == root && !d_rgb_mean) return fail(SPT_ERR_INVALID, "spt_render_gather: the root needs an output buffer");
  if (s->ctx->device != c->ctx->device) return fail(SPT_ERR_INVALID, "spt_render_gather: scene and communicator live on different devices");
  SptRenderDesc rd = *desc;
  rd.rank = c->rank; rd.world_size = c->world;
  int st = render_impl(s, &rd, nullptr, nullptr, stats);
  if (st != SPT_OK) return st;
  cudaStream_t cs = s->stream;
  const RenderParams& P = s->last_P;
  const int tw = P.tile_w, th = P.tile_h, tiles_x = P.tiles_x, tiles_y = (P.height + th - 1) / th, n_all = tiles_x * tiles_y;
  const size_t per_rank_tiles = (size_t)(n_all + c->world - 1) / c->world;            // ranks own n_all / world tiles, +-1
  const size_t tile_vals = (size_t)tw * th * 3, seg = per_rank_tiles * tile_vals;      // int64 values per rank segment
  const bool is_root = c->rank == root;
  const size_t need = (is_root ? (size_t)c->world : 1) * seg * sizeof(long long);
  if (s->gather_bytes < need) { cudaFree(s->gather_buf); s->gather_buf = nullptr; s->gather_bytes = 0; CU(cudaMalloc(&s->gather_buf, need)); s->gather_bytes = need; }
  CU(cudaEventRecord(s->sev[0], cs));
  const int key[6] = {P.width, P.height, tw, th, root, c->world};
  if (is_root && c->world > 1 && (!s->all_tiles || memcmp(key, s->all_tiles_key, sizeof(key)) != 0)) { // every rank's tile list, rank-major
    std::vector<uint32_t> all((size_t)c->world * per_rank_tiles, 0xffffffffu);
    for (int r = 0; r < c->world; r++) {
      const std::vector<uint32_t> t = owned_tiles(P.width, P.height, tw, th, r, c->world);
      std::copy(t.begin(), t.end(), all.begin() + (size_t)r * per_rank_tiles);
    }
    if (s->all_tiles_cap < all.size()) { cudaFree(s->all_tiles); s->all_tiles = nullptr; CU(cudaMalloc(&s->all_tiles, all.size() * sizeof(uint32_t))); s->all_tiles_cap = all.size(); }
    CU(cudaMemcpyAsync(s->all_tiles, all.data(), all.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, cs));
    CU(cudaStreamSynchronize(cs)); // `all` goes out of scope
    memcpy(s->all_tiles_key, key, sizeof(key));
  }
  if (c->world > 1) {
    if (!is_root) spt::launch_pack_tiles(s->film, s->tiles, (int)s->n_tiles, P, s->gather_buf, grid_for(s->ctx, (long long)s->n_tiles * tw * th, 8), cs);
    CU(cudaGetLastError());
    NC(g_nccl.GroupStart());
    if (is_root) { for (int r = 0; r < c->world; r++) if (r != root) NC(g_nccl.Recv(s->gather_buf + (size_t)r * seg, seg, ncclInt64, r, c->comm, cs)); }
    else NC(g_nccl.Send(s->gather_buf, seg, ncclInt64, root, c->comm, cs));
    NC(g_nccl.GroupEnd());
    if (is_root) { // the other ranks' tiles into the root's film (its own are there already)
      spt::launch_unpack_tiles(s->film, s->all_tiles, (int)per_rank_tiles, c->world, root, P, s->gather_buf, grid_for(s->ctx, (long long)c->world * per_rank_tiles * tw * th, 8), cs);
      CU(cudaGetLastError());
    }
  }
  if (is_root) resolve_film(s, rd.filter_radius, d_rgb_mean, nullptr, cs);
  CU(cudaEventRecord(s->sev[1], cs));
  CU(cudaStreamSynchronize(cs));
  CU(cudaGetLastError());
  if (stats) { float ms = 0.f; cudaEventElapsedTime(&ms, s->sev[0], s->sev[1]); stats->seconds_stage[5] = ms * 1e-3; stats->kernel_launches += is_root ? 2 : 1; }
  return SPT_OK;
}

extern "C" int32_t spt_render(SptScene* s, const SptRenderDesc* r, float* out_rgb_mean, uint8_t* out_rgb8, SptStats* stats) {
  if (!s || !r || !out_rgb_mean) return fail(SPT_ERR_INVALID, "spt_render: null argument");
  if (r->width <= 0 || r->height <= 0) return fail(SPT_ERR_INVALID, "spt_render: bad width/height");
  CU(cudaSetDevice(s->ctx->device));
  const size_t n_px = (size_t)r->width * r->height;
  if (s->out_px < n_px) {
    cudaFree(s->out_mean); cudaFree(s->out_rgb8); s->out_mean = nullptr; s->out_rgb8 = nullptr; s->out_px = 0;
    CU(cudaMalloc(&s->out_mean, n_px * 3 * sizeof(float))); CU(cudaMalloc(&s->out_rgb8, n_px * 3));
    s->out_px = n_px;
  }
  const int st = render_impl(s, r, s->out_mean, out_rgb8 ? s->out_rgb8 : nullptr, stats);
  if (st != SPT_OK) return st;
  CU(cudaMemcpyAsync(out_rgb_mean, s->out_mean, n_px * 3 * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
  if (out_rgb8) CU(cudaMemcpyAsync(out_rgb8, s->out_rgb8, n_px * 3, cudaMemcpyDeviceToHost, s->stream));
  CU(cudaStreamSynchronize(s->stream));
  return SPT_OK;
}

extern "C" int32_t spt_owned_tiles(int32_t width, int32_t height, int32_t tile_w, int32_t tile_h, int32_t rank, int32_t world_size,
                                   uint32_t* out, int64_t cap, int64_t* n_out) {
  const int tw = tile_w > 0 ? tile_w : SPT_DEFAULT_TILE_W, th = tile_h > 0 ? tile_h : SPT_DEFAULT_TILE_H;
  if (width <= 0 || height <= 0 || world_size <= 0 || rank < 0 || rank >= world_size || !n_out) return fail(SPT_ERR_INVALID, "spt_owned_tiles: bad argument");
  const std::vector<uint32_t> mine = owned_tiles(width, height, tw, th, rank, world_size);
  *n_out = (int64_t)mine.size();
  for (int64_t i = 0; i < (int64_t)mine.size() && i < cap && out; i++) out[i] = mine[i];
  return SPT_OK;
}
