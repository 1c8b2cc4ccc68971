/*
 * pt_oracle.h -- CPU restatement of the reference's render hot path (plain C, scalar).
 *
 * TEST INFRASTRUCTURE ONLY.  This is the checker, never the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * liboracle.  libpt_b200.so does not link, call or fall back to anything in this directory.
 *
 * Parity pin: with PTO_RNG_MT19937 the oracle consumes the same mt19937 stream, in the same
 * order, with the same float arithmetic (no FMA contraction) as the reference, and is checked
 * BIT-EXACT against oracle/_ref (the reference's own sources compiled here) in
 * tests/test_oracle_vs_reference.py, and against the committed fixtures in tests/golden/.
 * With PTO_RNG_PHILOX it draws the counter-based streams the CUDA kernels use (DESIGN.md,
 * "RNG streams"), so GPU paths can be compared with it sample by sample.
 */
#ifndef PT_ORACLE_H
#define PT_ORACLE_H

#include "pt_api.h"

#ifdef __cplusplus
extern "C" {
#endif

enum { PTO_RNG_MT19937 = 0, PTO_RNG_PHILOX = 1 };

/* seed / reseed the process-wide mt19937 used by PTO_RNG_MT19937 calls (ray_math.hpp:245-259) */
void pto_seed_mt(uint32_t seed);
/* n raw uniforms from that stream (== the reference's RandomUnilateral) */
void pto_random_mt(uint32_t n, float *out);
/* uniform k (k = 0,1,2,...) of Philox stream (seed; a, b, tag) */
float pto_philox_uniform(uint64_t seed, uint32_t a, uint32_t b, uint32_t tag, uint32_t k);

/* RayCastIntersect (main.cpp:406-556); triangles by brute force in index order */
void pto_intersect(const pt_scene *s, uint32_t n, const float *origins, const float *dirs, pt_hit *hits);

/* RayCast (main.cpp:558-823) for n rays.
   MT19937: sequential draws from the global stream.  PHILOX: path k uses stream (seed; k, 0, tag 2). */
void pto_raycast(const pt_scene *s, int rng_mode, uint64_t seed, uint32_t n, const float *origins,
                 const float *dirs, float *rgb);

/* RenderTexel's sampling loops (main.cpp:990-1186) over the window [x0,x0+w) x [y0,y0+h) of a
   width x height frame, samples [s_begin, s_end) of the pp*pp list (s_end = 0: all).
   Outputs (any may be NULL): sum / sumsq per pixel and channel in double over the accepted samples;
   mean_f = the reference's own float accumulation (color += contrib*radiance); ldr = packed BGRA.
   MT19937 mode requires s_begin = 0, s_end = pp*pp (the stream is sequential). */
void pto_render(const pt_scene *s, int width, int height, int pp, int rng_mode, uint64_t seed,
                int x0, int y0, int w, int h, uint32_t s_begin, uint32_t s_end,
                double *sum, double *sumsq, float *mean_f, uint32_t *ldr);

/* same, PHILOX only, rows split over `threads` pthreads; returns seconds spent rendering */
double pto_render_threads(const pt_scene *s, int width, int height, int pp, uint64_t seed, int threads,
                          float *mean_f, uint32_t *ldr);

/* TonemapPass + LinearToSRGB + BGRAPack4x8 (main.cpp:1664-1673, 1175-1182; ray_math.hpp:196-202, 320-332) */
void pto_resolve(uint32_t n, const float *rgb, uint32_t *bgra);

/* BespokeSampleTexture (main.cpp:1675-1709) on scene texture idx (0..3) */
void pto_sample_texture(const pt_scene *s, int idx, float u, float v, float *rgb);

/* jitter-free ray through the lower-left corner of every pixel (main.cpp:1003-1006, 1059-1064) */
void pto_primary_rays(const pt_scene *s, int width, int height, float *origins, float *dirs);

#ifdef __cplusplus
}
#endif
#endif
