/*
 * pt_api.h -- C-ABI of the B200 path-tracing render core (libpt_b200.so).
 *
 * Drop-in boundary.  The reference (BluBloos/Pathtracer) has no FFI; its render seam is
 *     void RenderTexel(unsigned int *out, texel_t texel)        src/win32_main.cpp:45, 990-1186
 * called per 32x32 tile by RenderThread (src/win32_main.cpp:851-875) for tiles issued by
 * MasterThread (src/win32_main.cpp:879-987), reading file-static globals g_world, g_camera,
 * g_textures[4], g_pp, g_worldKind, g_bNormals/g_bMetalness/g_bRoughness
 * (src/win32_main.cpp:97-118).  This library replaces the body of MasterThread between the
 * tile-list creation and WriteDIBImage (src/win32_main.cpp:888-984):
 *
 *     pt_upload_scene  <- what LoadWorld/DefineCamera left in the globals (main.cpp:233-236)
 *     pt_render        <- the RenderThread/RenderTexel fan-out           (main.cpp:906-978)
 *     pt_readback      <- g_image.pixelPointer, row 0 = bottom           (main.cpp:863-866, 1181-1182)
 *
 * Plain pointers and sizes only; no C++/torch types.  All structs are packed to 4 bytes,
 * little-endian, and mirror the reference PODs of include/ray.hpp field for field so a
 * maintainer can hand the reference's own arrays over without conversion.
 * Every entry point returns 0 on success or a negative pt_status; the message is available
 * through pt_last_error().  There is NO CPU fallback: without a CUDA device pt_create fails.
 */
#ifndef PT_API_H
#define PT_API_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PT_API_VERSION 1

typedef enum pt_status {
    PT_OK = 0,
    PT_ERR_INVALID_ARG = -1,
    PT_ERR_NO_DEVICE = -2,      /* no usable CUDA device: the library never falls back to the CPU */
    PT_ERR_CUDA = -3,
    PT_ERR_NO_SCENE = -4,
    PT_ERR_NO_FRAME = -5,
    PT_ERR_OOM = -6,
    PT_ERR_UNSUPPORTED = -7
} pt_status;

#pragma pack(push, 4)

typedef struct pt_v3 { float x, y, z; } pt_v3;

/* material_t, include/ray.hpp:36-79 (68 bytes).  Texture indices are 1-based, 0 = none. */
typedef struct pt_material {
    float   alpha;            /* unused by the hot path (ray.hpp:63) */
    float   ior;              /* F0 = ((1.003-ior)/(1.003+ior))^2, main.cpp:600 */
    int32_t albedo_tex;
    pt_v3   albedo;
    int32_t metalness_tex;
    float   metalness;
    pt_v3   metal_color;
    int32_t roughness_tex;
    float   roughness;
    int32_t normal_tex;
    pt_v3   emit;             /* non-zero => light; material 0 is the sky (main.cpp:579-581, 825-827) */
} pt_material;

typedef struct pt_plane  { pt_v3 n; float d; uint32_t mat; } pt_plane;                  /* ray.hpp:83-87  (20 B) */
typedef struct pt_quad   { pt_v3 point, u, v; uint32_t mat; } pt_quad;                  /* ray.hpp:89-94  (40 B) */
typedef struct pt_sphere { pt_v3 p; float r; uint32_t mat; } pt_sphere;                 /* ray.hpp:96-100 (20 B) */
typedef struct pt_aabb   { pt_v3 origin, half_dim, min, max; uint32_t mat; } pt_aabb;   /* ray_math.hpp:136-144 (52 B) */

/* camera_t, include/ray.hpp:176-186 (116 bytes), as DefineCamera leaves it (main.cpp:2197-2231) */
typedef struct pt_camera {
    float   fov;
    float   focal_length;
    float   focal_distance;
    float   aperture_radius;
    uint8_t use_pinhole;
    uint8_t _pad[3];
    float   film_w, film_h, half_film_w, half_film_h, half_pixel_w, half_pixel_h;
    pt_v3   pos, target, frustum_center, axis_z, axis_x, axis_y;
} pt_camera;

/* texture_t mip 0, include/ray.hpp:143-146.  Texels cross the boundary as the RGBA8 bytes
   stb_image produced; byte/255.f on fetch is bit-identical to main.cpp:1737-1739. */
typedef struct pt_texture {
    int32_t        width, height;
    const uint8_t *rgba8;     /* width*height*4 bytes, row-major, R,G,B,A; NULL = absent */
} pt_texture;

enum {
    PT_SCENE_NORMAL_MAPS    = 1u << 0,   /* g_bNormals    (cleared by -n, main.cpp:2173) */
    PT_SCENE_METALNESS_MAPS = 1u << 1,   /* g_bMetalness  (cleared by -m, main.cpp:2175) */
    PT_SCENE_ROUGHNESS_MAPS = 1u << 2,   /* g_bRoughness  (cleared by -r, main.cpp:2177) */
    PT_SCENE_JUST_COSINE    = 1u << 3,   /* g_worldKind == WORLD_RAYTRACING_ONE_WEEKEND (main.cpp:654-655) */
    /* Upload hint, no reference counterpart (LoadTexture runs once, main.cpp:1711-1746): a texture whose (rgba8 pointer,
       width, height) equal those of the previous pt_upload_scene on this context has unchanged texels; the device copy is
       kept and nothing is sent.  Without the flag every upload re-sends every texture. */
    PT_SCENE_KEEP_TEXTURES  = 1u << 8
};

/* world_t + camera + textures + flags: everything RenderTexel reads (ray.hpp:152-162) */
typedef struct pt_scene {
    uint32_t           struct_size;      /* = sizeof(pt_scene) */
    uint32_t           flags;            /* PT_SCENE_* */
    uint32_t           n_materials;
    uint32_t           n_spheres;        /* spheres[0] is the importance-sampled light (main.cpp:683) */
    uint32_t           n_planes;
    uint32_t           n_quads;
    uint32_t           n_aabbs;
    uint32_t           n_triangles;
    const pt_material *materials;
    const pt_sphere   *spheres;
    const pt_plane    *planes;
    const pt_quad     *quads;
    const pt_aabb     *aabbs;
    const float       *tri_points;       /* n_triangles * 9 floats: A, B, C (mesh_t.points, ray.hpp:102-106) */
    const int32_t     *tri_materials;    /* one per triangle (= mesh_t.matIndices[3*tri], main.cpp:520) */
    pt_texture         textures[4];      /* g_textures[0..3]: albedo, metallic, roughness, normal (main.cpp:1720-1723) */
    pt_camera          camera;
} pt_scene;

/* BVH builder (replaces GenerateAccelerationStructure, main.cpp:1188-1452); both give bit-identical images */
enum { PT_BVH_HOST_SAH = 0,       /* surface-area heuristic on the host (default; best traversal) */
       PT_BVH_DEVICE_LBVH = 1 };  /* Morton codes -> radix sort -> Karras hierarchy -> refit, on the GPU */
/* render kernel; both run the same samples (Philox streams keyed by pixel and sample) */
enum { PT_KERNEL_AUTO = 0,        /* = PT_KERNEL_MEGA */
       PT_KERNEL_MEGA = 1,        /* path-regeneration megakernel, feature-specialised variants (default); on a triangle tree
                                     the walks that few lanes of a warp enter are traced one ray at a time by all 32 lanes
                                     over a 32-cluster cut of the BVH -- same hits, bit-identical images */
       PT_KERNEL_WAVEFRONT = 2 }; /* paths pooled in shared memory, one index list per stage, per warp */

typedef struct pt_config {
    uint32_t struct_size;     /* = sizeof(pt_config) */
    int32_t  device;          /* CUDA ordinal */
    int32_t  bvh_builder;     /* PT_BVH_* */
    int32_t  kernel;          /* PT_KERNEL_* */
    void    *stream;          /* cudaStream_t to launch on; NULL = a stream the context owns */
} pt_config;

/* Scheduling knobs of the kernels and the BVH builder.  None of them changes a sample: images are bit-identical for
 * every setting (tests/test_gpu_properties.py); they exist so that a host -- and those tests -- can pin a schedule.
 * A field left at its "default" value keeps the measured default.  The library reads no environment variable. */
typedef struct pt_tuning {
    uint32_t struct_size;        /* = sizeof(pt_tuning) */
    int32_t  sample_lanes_log2;  /* lanes of a warp that share a pixel = 1 << n, n in 0..5; -1 = default (by launch size) */
    int32_t  shade_batch;        /* parked lanes before the BRDF code runs, 1..32; 0 = default */
    int32_t  coop_max;           /* walks entered by <= n lanes are traced by the whole warp, 0..32 (0 = never); -1 = default.
                                    The cluster cut is built by pt_upload_scene: set this BEFORE uploading the scene. */
    int32_t  tri_leaf, sph_leaf; /* max primitives per BVH leaf, 1..64; 0 = default (4 / 3).  Applies to the next upload. */
    int32_t  sah_sweep;          /* 1 = exact sweep for small nodes (default), 0 = 16-bin SAH everywhere; -1 = default */
    int32_t  wf_pixels_log2;     /* PT_KERNEL_WAVEFRONT: pixels per warp pool; 0 = default */
} pt_tuning;

enum {
    PT_RENDER_CLEAR = 1u << 0     /* zero the accumulation buffer before rendering */
};

/*
 * One pt_render call = samples [sample_begin, sample_end) of the p*p sample list of every pixel of
 * every 32x32 tile (the reference's tile grid, main.cpp:888-904) that belongs to tile_rank,
 * ADDED into the float4 accumulation buffer.  Tile (tx, ty) belongs to rank (tx + k * ty) % tile_count with k coprime
 * to tile_count (pt_tile_owner): a diagonal interleave, so that no rank gets a vertical stripe of the image whatever the
 * frame width.  Which rank renders a tile never changes its pixels (the RNG is keyed by pixel and sample).  pp as g_pp (main.cpp:113): spp = pp*pp.
 * sample index s = i*pp + j for the pinhole strata loop (main.cpp:1041-1042) and
 * s = rayIndex*pp + rayIndex2 for the lens loops (main.cpp:1114, 1145).
 * sample_end == 0 means pp*pp.  tile_count == 0 means 1.
 */
typedef struct pt_render_params {
    uint32_t struct_size;     /* = sizeof(pt_render_params) */
    int32_t  width, height;
    int32_t  pp;
    uint64_t seed;            /* Philox key */
    uint32_t sample_begin, sample_end;
    uint32_t tile_rank, tile_count;
    uint32_t flags;           /* PT_RENDER_* */
} pt_render_params;

typedef struct pt_stats {
    uint64_t samples;         /* camera paths accumulated so far (all pt_render calls since the last clear) */
    uint64_t rays;            /* closest-hit queries, if counted (debug builds), else 0 */
    uint64_t nan_retries;     /* samples redrawn because the radiance was NaN (main.cpp:1068, 1166) */
    float    render_ms;       /* device time of the last pt_render (CUDA events on the launch stream) */
    uint32_t kernel_launches; /* kernels launched by the last pt_render */
    uint32_t bvh_nodes;
    uint32_t bvh_prims;
    uint64_t nan_dropped;     /* samples still NaN after 8 redraws, accumulated as black (the reference would redraw for ever) */
    uint64_t texture_bytes_uploaded; /* texel bytes the last pt_upload_scene actually sent (0 = all textures were unchanged) */
} pt_stats;

/* primitive id of a closest hit; kind order = the reference's test order (main.cpp:418-553) */
enum { PT_HIT_MISS = 0, PT_HIT_SPHERE = 1, PT_HIT_QUAD = 2, PT_HIT_PLANE = 3, PT_HIT_TRIANGLE = 4, PT_HIT_AABB = 5 };
typedef struct pt_hit {
    float    t;               /* FLT_MAX on miss (main.cpp:411) */
    int32_t  kind;            /* PT_HIT_* */
    int32_t  index;           /* index within its primitive array; -1 on miss */
    uint32_t mat;             /* hitMatIndex; 0 on miss (main.cpp:412) */
    pt_v3    normal;          /* ray_payload_t.normal (ray.hpp:137-141) */
} pt_hit;

#pragma pack(pop)

typedef struct pt_context pt_context;

/* lifetime */
int         pt_create(const pt_config *cfg, pt_context **out);
int         pt_destroy(pt_context *ctx);
const char *pt_last_error(const pt_context *ctx);      /* ctx may be NULL: error of the last failed pt_create */
int         pt_api_version(void);

/* scene: copies everything; the caller's pointers need not outlive the call.  Builds the BVH. */
int pt_upload_scene(pt_context *ctx, const pt_scene *scene);
/* replace only the camera (e.g. a new resolution) without rebuilding the BVH */
int pt_set_camera(pt_context *ctx, const pt_camera *cam);
/* scheduling knobs (see pt_tuning); takes effect at the next pt_upload_scene / pt_render */
int pt_set_tuning(pt_context *ctx, const pt_tuning *tuning);

/* render: asynchronous on the context's stream; pt_readback / pt_synchronize wait for it */
int pt_render(pt_context *ctx, const pt_render_params *params);
int pt_synchronize(pt_context *ctx);

/* resolve (accum/spp -> ACES -> sRGB -> BGRA8, main.cpp:1172-1182) on the device */
int pt_resolve(pt_context *ctx);

/* host readback.  bgra: width*height uint32 (0xAARRGGBB, row 0 = bottom), may be NULL.
   accum_rgba: width*height*4 floats (sum of radiance r,g,b and the sample count), may be NULL.
   Runs pt_resolve if bgra != NULL. */
int pt_readback(pt_context *ctx, uint32_t *bgra, float *accum_rgba, pt_stats *stats);

/* device-side access for multi-GPU plumbing (NCCL reduce / gather run on these buffers) */
int pt_device_buffers(pt_context *ctx, void **accum_rgba_dev, void **bgra_dev, int32_t *width, int32_t *height);
int pt_set_stream(pt_context *ctx, void *cuda_stream);
/* Tile -> rank map: the owner of tile (tx, ty) of a width x height frame split over tile_count ranks, and (slot_out, may be
 * NULL) the tile's slot in that rank's pack.  pt_tile_slots = slots per rank (the same for every rank; >= its tile count:
 * a slot whose tile falls outside the grid is empty and packs as zeros). */
uint32_t pt_tile_owner(int32_t width, int32_t height, uint32_t tile_count, uint32_t tx, uint32_t ty, uint32_t *slot_out);
uint32_t pt_tile_slots(int32_t width, int32_t height, uint32_t tile_count);
/* pack this rank's tile slots (slot-major, 32*32 uint32 each, zero-padded) / scatter a packed set back; *n_tiles_out = slots */
int pt_pack_tiles(pt_context *ctx, void *dst_dev, uint32_t tile_rank, uint32_t tile_count, uint32_t *n_tiles_out);
int pt_unpack_tiles(pt_context *ctx, const void *src_dev, uint32_t tile_rank, uint32_t tile_count);
/* The same exchange in two kernels per frame instead of resolve + pack + (tile_count - 1) unpacks:
 *   pt_resolve_pack_tiles : resolve THIS rank's tiles from the accumulator straight into the tile-major pack;
 *   pt_unpack_gathered    : scatter the packs of all tile_count ranks, gathered back to back with stride_words uint32
 *                           per rank, into this context's BGRA frame (one thread per pixel) and mark it resolved. */
int pt_resolve_pack_tiles(pt_context *ctx, void *dst_dev, uint32_t tile_rank, uint32_t tile_count, uint32_t *n_tiles_out);
int pt_unpack_gathered(pt_context *ctx, const void *src_dev, uint32_t tile_count, uint32_t stride_words);
/* Sample-range sharding with a reduce-scatter: resolve n_pixels of ANY device accumulator (float4 per pixel: a rank's
 * share of the reduced sums) into ANY device BGRA buffer; pt_mark_resolved tells the context that its BGRA frame was
 * filled by the caller (gathered shares), so that pt_readback does not resolve again. */
int pt_resolve_range(pt_context *ctx, const void *accum_dev, void *bgra_dev, uint32_t n_pixels);
int pt_mark_resolved(pt_context *ctx);

/* Single-process multi-GPU (one context per device, same frame size): exchange over NVLink peer copies,
 * ordered by events on the two contexts' streams -- the counterpart of the gather / reduce that
 * torch.distributed does for one process per GPU.  No reference counterpart (CPU threads share g_image,
 * main.cpp:107).
 *   pt_gather_tiles_from: resolve + pack the tiles of src that belong to tile_rank, copy them to dst's
 *                         device and unpack them into dst's BGRA frame.
 *   pt_add_accum_from:    dst.accum += src.accum (sample-range sharding); dst must be resolved again. */
int pt_gather_tiles_from(pt_context *dst, pt_context *src, uint32_t tile_rank, uint32_t tile_count);
int pt_add_accum_from(pt_context *dst, pt_context *src);

/* One process per GPU, direct tile exchange over NVLink peer memory (the fused "resolve + gather"): the root rank exports
 * its BGRA frame as a CUDA IPC handle, every other rank maps it, and pt_resolve_tiles_to resolves THIS rank's tiles
 * (pt_tile_owner == tile_rank) and stores the BGRA words straight into the mapped frame (one 128-byte row segment of a
 * tile per warp) -- no pack, no NCCL gather, no unpack.  The caller orders the ranks (a stream-ordered all-reduce of one
 * word before and after; pathtracer_b200/dist.py, mode "tiles_p2p").  No reference counterpart: the reference's worker
 * threads write their tiles into the one shared g_image (main.cpp:107, 862-866) -- this is that, across GPUs.
 *   pt_frame_ipc_handle : allocate the width x height frame if needed; write the PT_IPC_HANDLE_BYTES blob of its BGRA buffer.
 *                         An exported frame keeps its size: a later pt_render at another size fails with PT_ERR_INVALID_ARG.
 *                         Fails with PT_ERR_CUDA when the driver cannot name the frame's allocation (nothing is exported).
 *   pt_open_peer_frame  : map another process's frame; *remote_bgra is a device pointer valid in this process.
 *   pt_resolve_tiles_to : dst_bgra = NULL resolves into the context's own frame; otherwise dst_bgra must be a frame this
 *                         context mapped with pt_open_peer_frame, of the same width x height (checked). */
#define PT_IPC_HANDLE_BYTES 80        /* cudaIpcMemHandle_t (64) + the frame's byte offset inside the exported allocation (8) + width, height (4 + 4) */
int pt_frame_ipc_handle(pt_context *ctx, int32_t width, int32_t height, void *handle_out);
int pt_open_peer_frame(pt_context *ctx, const void *handle, void **remote_bgra);
int pt_close_peer_frame(pt_context *ctx, void *remote_bgra);
int pt_resolve_tiles_to(pt_context *ctx, void *dst_bgra, uint32_t tile_rank, uint32_t tile_count);

/* debug / parity exports */
/* closest hit of n caller-supplied rays (origins, dirs: n*3 floats, host pointers) */
int pt_trace_rays(pt_context *ctx, uint32_t n, const float *origins, const float *dirs, pt_hit *hits_out);
/* closest hit of the jitter-free ray through the lower-left corner of every pixel (row 0 = bottom) */
int pt_primary_hits(pt_context *ctx, int32_t width, int32_t height, pt_hit *hits_out);
/* radiance of n caller-supplied camera rays; path k draws from Philox stream (seed, k, 0) */
int pt_trace_paths(pt_context *ctx, uint32_t n, const float *origins, const float *dirs, uint64_t seed, float *rgb_out);
/* resolve n HDR colours (n*3 floats, host) to BGRA8 exactly as pt_resolve does */
int pt_resolve_colors(pt_context *ctx, uint32_t n, const float *rgb, uint32_t *bgra_out);
/* The same as pt_resolve_colors for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  NO COUNTERPART IN THE REFERENCE.  It allocates nothing -- no
 * cudaMalloc / cudaFree, no allocating library routine: the staging that pt_resolve_colors allocates per call lives in a fixed
 * 1 MiB device array inside the library (held by every GPU that loads the library until the process exits), so
 * n <= PT_NOALLOC_MAX_COLORS; a larger n is rejected with PT_ERR_UNSUPPORTED and a message (split the call, or use
 * pt_resolve_colors).  rgb and bgra_out are HOST pointers supplied by the caller, as for pt_resolve_colors; results are
 * byte-identical to it.  The scratch area is shared by every call on the same GPU and keeps the last call's contents (nothing
 * in it needs to start at zero: each call overwrites what it reads); calls that overlap on the GPU -- contexts on different
 * streams with no ordering between them -- race on it.  The call returns after its own work has finished. */
#define PT_NOALLOC_MAX_COLORS 65536u
int pt_resolve_colors_noalloc(pt_context *ctx, uint32_t n, const float *rgb, uint32_t *bgra_out);

/* measurement helper for the roofline: achieved FP32 FMA throughput of this device (TFLOP/s, best of 5
   runs of an 8-way-ILP FMA chain on every SM), timed with CUDA events on the context's stream */
int pt_measure_fp32_peak(pt_context *ctx, float *tflops_out);
/* measurement helper: memset a 256 MiB scratch buffer (> the 126 MB L2) on the context's stream */
int pt_flush_l2(pt_context *ctx);

#ifdef __cplusplus
}
#endif
#endif /* PT_API_H */
