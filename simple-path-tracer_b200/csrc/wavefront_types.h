// wavefront_types.h — plain structs shared by the host render loop (spt_api.cu) and the wavefront kernels.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

#define SPT_NQ 10        // shade queues: 0..7 = material tag, 8 = special (medium / environment), 9 = unsorted surfaces
#define SPT_Q_SPECIAL 8
#define SPT_Q_GENERIC 9

// Per-path state written by shade / generate: ONE 64-byte record per slot (two full 32-byte sectors), because the shade
// kernels visit slots in queue order (a gather): with separate arrays every 16-byte field cost its own half-used sector.
struct __align__(16) PathRec {
  float4 ray_o;   // origin.xyz, t_min            } sector 0: all that k_extend reads
  float4 ray_d;   // direction.xyz, -             }
  float4 thr;     // throughput.rgb, last_sample_pdf
  uint4 aux;      // rng state (lo, hi), pixel index j*W+i, sample index k within the pixel
};
struct PathPool { // `n` slots
  PathRec* rec;
  float4* hit;      // t, u, v, global triangle index (int bits; -1 = sphere)   written by k_extend
  int32_t* hit_inst;
  uint32_t* flags;  // bit0 alive | depth << 8 | (medium+1) << 16  (dense passes read only this)
  int32_t n;
};
struct ShadowQueue { float4* o; float4* d; float4* c; }; // o.xyz,tmin | d.xyz,tmax | contrib.rgb,pixel
struct IterCounters { // per-iteration device counters; each is zeroed by a kernel that runs between its consumer and its next producer
  uint32_t q[SPT_NQ];   // shade-queue fill      (producer: logic/special, consumers: special/shade, zeroed by generate)
  uint32_t sq;          // shadow-queue fill     (producer: special/shade, consumer: connect, zeroed by logic)
  uint32_t live;        // live slots after generate (read by the host; zeroed by logic)
  uint32_t next_ext;    // ray-fetch cursor of extend  (zeroed by logic)
  uint32_t next_con;    // ray-fetch cursor of connect (zeroed by logic)
  uint32_t drain;       // 1: every camera sample has been started (set by logic): generate compacts the live slots into a list and extend
                        //    fetches from that list instead of scanning a mostly empty pool
  uint32_t n_ext[2];    // entries of ext_queue[parity] (producer: generate in drain mode, consumers: extend of the same iteration, logic and
                        //    generate of the NEXT one; n_ext[parity] is zeroed by logic of its own iteration)
  uint32_t list_valid[2]; // ext_queue[parity] lists the live slots left by that iteration (written by generate only: a kernel never reads
                        //    the flag of its own parity, so it is stable while it runs)
};
struct DevCounters { // device-side totals of one render
  unsigned long long cursor;      // next camera-sample ordinal
  unsigned long long paths, ext_rays, shadow_rays, probe_rays, nonfinite, nodes, tris, insts, shadow_culled;
  unsigned long long trav_profile[10]; // COUNT builds: see TravCounters
  IterCounters it;
};
struct RenderParams {
  PathPool pool;
  ShadowQueue sq;
  uint32_t* queue;   // SPT_NQ arrays of pool.n slot ids
  uint32_t* ext_queue[2]; // pool.n slot ids each: the live slots of a draining pool (see IterCounters), one list per iteration parity
  int32_t parity;         // iteration index & 1
  DevCounters* ctr;
  long long* film; // 3 per pixel
  const uint32_t* tiles; // owned tile ids
  int32_t n_tiles, tiles_x, tile_w, tile_h;
  int32_t width, height, max_depth;
  int32_t sampler, spp, div_x, div_y; // spp = samples per pixel rendered by THIS call
  int32_t spp_first, spp_frame;       // index of its first sample; samples per pixel of the whole frame (sampler ordinals)
  int32_t sort_materials; // 0: every surface hit goes to the generic queue (SPT_RENDER_NO_SORT)
  int32_t use_ext_queue;  // 1: a draining pool is compacted for extend (IterCounters::drain); SPT_EXT_QUEUE=0 turns it off for A/B timing
  uint64_t seed;
  unsigned long long total_samples; // n_tiles * tile_w * tile_h * spp
};
