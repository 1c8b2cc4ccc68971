/*
 * pt_host.h -- host side of the drop-in (libpt_host.so): everything the reference does on the CPU
 * before and after the render seam, restated from scratch so the `pathtracer` executable keeps the
 * reference's surface (flags -t/-p/-w/-d/-n/-m/-r/-h, the five worlds, GLB import, .BMP output).
 *
 *   pt_host_parse_args     <- ParseArgs            src/win32_main.cpp:2110-2195
 *   pt_host_load_world     <- LoadWorld            src/win32_main.cpp:1788-2046 (+ LoadGltf 1454-1593,
 *                                                  LoadBespokeTextures/LoadTexture 1711-1744)
 *   pt_host_define_camera  <- DefineCamera         src/win32_main.cpp:2197-2231
 *   pt_host_write_bmp      <- WriteDIBImage        src/win32_main.cpp:358-391
 *
 * No render arithmetic lives here: pixels come only from libpt_b200.so (include/pt_api.h).
 */
#ifndef PT_HOST_H
#define PT_HOST_H

#include "pt_api.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pt_host_world pt_host_world;

/* options exactly as ParseArgs leaves its globals (main.cpp:2111-2119), plus the new long options */
typedef struct pt_host_options {
    int32_t threads;          /* -t  g_tc: CPU threads; accepted, used only by the CPU baseline */
    int32_t pp;               /* -p  g_pp, clamped to [0,1000] (main.cpp:2171) */
    int32_t world;            /* -w  1..5, clamped (main.cpp:2181) */
    int32_t lens;             /* -d  thin lens on (main.cpp:2183) */
    int32_t normals;          /* -n clears (main.cpp:2173) */
    int32_t metalness;        /* -m clears (main.cpp:2175) */
    int32_t roughness;        /* -r clears (main.cpp:2177) */
    int32_t help;             /* -h */
    /* new, non-conflicting long options */
    int32_t width, height;    /* --width/--height (reference: fixed 1280x720 window, main.cpp:218-219) */
    int32_t gpus;             /* --gpus N */
    int32_t shard_samples;    /* --shard samples (default tiles) */
    uint64_t seed;            /* --seed: Philox key of the render */
    uint32_t scene_seed;      /* --scene-seed: mt19937 seed for world 4's layout (reference: random_device) */
    int32_t lbvh;             /* --lbvh: device LBVH instead of host SAH */
    int32_t kernel;           /* --kernel mega|wavefront */
    char    out_path[512];    /* --out (reference: "test.bmp", main.cpp:984) */
    char    res_dir[512];     /* --res (reference: "res\\", main.cpp:1459) */
    int32_t system_cores;     /* what "System has %d core(s)." printed */
    int32_t snapshots;        /* --snapshots N: write the .BMP after each of N sample chunks (the reference's live
                                 viewer, main.cpp:252-274, without a window) */
    int32_t hdr_dump;         /* --hdr: also write <out>.hdr.bin (W*H float4: sum r,g,b + sample count) for image_compare */
} pt_host_options;

/* Parses argv with the reference's per-character scan; prints the same warnings/messages to stdout.
   Returns 0, or 1 if -h was given (help printed, caller should exit 0). */
int  pt_host_parse_args(int argc, char **argv, pt_host_options *opt);
void pt_host_print_help(void);

/* Build world 1..5.  res_dir holds mario.glb and the rusty-metal_*.png files.  scene_seed seeds the
   mt19937 that lays out world 4.  use_lens/flags as ParseArgs set them. */
int  pt_host_load_world(int world, const char *res_dir, uint32_t scene_seed, pt_host_world **out);
void pt_host_free_world(pt_host_world *w);
const char *pt_host_last_error(void);

/* Fill a pt_scene (pointers into the world; valid until pt_host_free_world) for a W x H frame.
   lens = -d flag; tex_flags = PT_SCENE_*_MAPS bits that are enabled. */
int  pt_host_scene(pt_host_world *w, int width, int height, int lens, uint32_t tex_flags, pt_scene *out);

/* DefineCamera: fills the derived fields from pos/target/fov/focal_distance/aperture_radius/use_pinhole */
void pt_host_define_camera(pt_camera *cam, int width, int height);

/* 58-byte-header 32 bpp bottom-up BMP, byte-identical to WriteDIBImage's output */
int  pt_host_write_bmp(const char *path, const uint32_t *bgra, int width, int height);
int  pt_host_read_bmp(const char *path, uint32_t **bgra_out, int *width, int *height);   /* free() the result */

/* decoders used by the loaders, exposed for tests */
int  pt_host_load_png_rgba8(const char *path, uint8_t **rgba_out, int *width, int *height);   /* free() the result */
int  pt_host_load_glb(const char *path, float **tri_points_out, int32_t **tri_mats_out, uint32_t *n_tris,
                      pt_material **new_materials_out, uint32_t *n_new_materials, uint32_t first_new_index);

#ifdef __cplusplus
}
#endif
#endif
