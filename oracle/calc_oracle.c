/*
 * calc_oracle.c — CPU restatement of the reference's escape-time path. TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
 * library; the product (fractal_renderer_b200/) never does and has no CPU fallback.
 *
 * PARITY UNPINNED by reference tests: Icelk/fractal-renderer has no tests, golden vectors or known-answer
 * fixtures for this path, and its Rust toolchain is absent here, so the reference itself
 * cannot be run. What pins this file instead: (1) a by-hand orbit KAT (tests/test_oracle.py), (2) the
 * independent survey restatement's SHA-256 hashes and known-answer pixels, (3) the
 * reference's only artefact, screenshots/mandelbrot-1000000x.avif, matched at interior IoU >= 0.9999
 * (tests/golden/make_screenshot_fixture.py).
 *
 * Every function cites the reference lines it follows (paths relative to /root/reference). Arithmetic is
 * IEEE-754 binary64, one rounding per written operator, left to right; build with -ffp-contract=off
 * (oracle/Makefile) so that GCC never fuses a*b+c. sqrt/log2 are glibc's, which is what Rust std's
 * f64::sqrt / f64::log2 call on this platform.
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/frb.h"

typedef struct { double re, im; } Im;

/* Imaginary::square — calc/src/lib.rs:87-92: re = (re*re) - (im*im); im = 2.0 * re * im (left to right). */
static inline Im im_square(Im z)
{
    Im r;
    r.re = (z.re * z.re) - (z.im * z.im);
    r.im = (2.0 * z.re) * z.im;
    return r;
}

/* Imaginary::squared_distance — calc/src/lib.rs:94-96. */
static inline double im_squared_distance(Im z) { return z.re * z.re + z.im * z.im; }

/* impl Add for Imaginary — calc/src/lib.rs:98-107. */
static inline Im im_add(Im a, Im b)
{
    Im r;
    r.re = a.re + b.re;
    r.im = a.im + b.im;
    return r;
}

/* recursive — calc/src/lib.rs:245-257. Returns the count, writes the final position. */
uint32_t oracle_recursive(uint32_t iterations, double start_re, double start_im, double c_re, double c_im,
                          double limit, double *out_re, double *out_im)
{
    const double squared = limit * limit;
    Im previous = { start_re, start_im };
    const Im c = { c_re, c_im };
    for (uint32_t i = 0; i < iterations; i++) {
        Im next = im_add(im_square(previous), c);
        double dist = im_squared_distance(next);
        if (dist > squared) {
            *out_re = next.re;
            *out_im = next.im;
            return i;
        }
        previous = next;
    }
    *out_re = previous.re;
    *out_im = previous.im;
    return iterations;
}

/* coord_to_space — calc/src/lib.rs:181-184. */
static inline double coord_to_space(double coord, double max, double offset, double pos, double scale)
{
    return ((coord / max) - offset) / scale + pos;
}

/* xy_to_imaginary — calc/src/lib.rs:185-197 (x is divided by the HEIGHT). */
void oracle_xy_to_imaginary(const FrbConfig *cfg, uint32_t x, uint32_t y, double *re, double *im)
{
    const double width = (double)cfg->width, height = (double)cfg->height;
    *re = coord_to_space((double)x, height, (width / height) / 2.0, cfg->pos.re, cfg->scale.re);
    *im = coord_to_space((double)y, height, 0.5, cfg->pos.im, cfg->scale.im);
}

/* Rust `f64 as u8`: truncate toward zero, saturate to [0,255], NaN -> 0. */
static inline uint8_t sat_u8(double v)
{
    if (!(v == v)) return 0;
    if (v <= 0.0) return 0;
    if (v >= 255.0) return 255;
    return (uint8_t)v;
}

/* RGB::new(r, b, g) — calc/src/lib.rs:129-131: second argument is stored in .b, third in .g. */
static inline FrbRGB rgb_new(uint8_t r, uint8_t b, uint8_t g)
{
    FrbRGB c;
    c.r = r;
    c.g = g;
    c.b = b;
    return c;
}

/* color_multiply — calc/src/lib.rs:133-139: RGB::new(color.r*m, color.g*m, color.b*m) => g and b swap. */
static inline FrbRGB color_multiply(FrbRGB color, double mult)
{
    return rgb_new(sat_u8((double)color.r * mult), sat_u8((double)color.g * mult), sat_u8((double)color.b * mult));
}

/* The colour half of get_recursive_pixel — calc/src/lib.rs:214-234, given recursive()'s result. */
FrbRGB oracle_shade(const FrbConfig *cfg, double z_re, double z_im, uint32_t iters_u)
{
    Im pos = { z_re, z_im };
    double dist = im_squared_distance(pos);
    if (dist > cfg->stable_limit) {
        double iters = (double)iters_u;
        if (cfg->smooth) {
            double log_zn = log2(sqrt(dist)) / 2.0;
            double nu = log2(log_zn);
            iters += 1.0 - nu;
        }
        double mult = iters / (double)cfg->iterations * cfg->exposure;
        return color_multiply(cfg->primary_color, mult);
    } else if (cfg->inside) {
        return color_multiply(cfg->secondary_color, dist);
    }
    return rgb_new(0, 0, 0);
}

/* get_recursive_pixel — calc/src/lib.rs:199-235. Optionally reports recursive()'s raw result. */
FrbRGB oracle_get_recursive_pixel_raw(const FrbConfig *cfg, uint32_t x, uint32_t y, uint32_t *iters_out,
                                      double *z_out /* 2 doubles or NULL */)
{
    double sre, sim, zre, zim;
    uint32_t iters;
    oracle_xy_to_imaginary(cfg, x, y, &sre, &sim);
    switch (cfg->algo) {
    case FRB_ALGO_MANDELBROT:
        iters = oracle_recursive(cfg->iterations, sre, sim, sre, sim, cfg->limit, &zre, &zim);
        break;
    case FRB_ALGO_JULIA:
        iters = oracle_recursive(cfg->iterations, sre, sim, cfg->julia_set.re, cfg->julia_set.im, cfg->limit,
                                 &zre, &zim);
        break;
    default: /* `_ => return RGB::BLACK` (:211) */
        if (iters_out) *iters_out = 0;
        if (z_out) { z_out[0] = 0.0; z_out[1] = 0.0; }
        return rgb_new(0, 0, 0);
    }
    if (iters_out) *iters_out = iters;
    if (z_out) { z_out[0] = zre; z_out[1] = zim; }
    return oracle_shade(cfg, zre, zim, iters);
}

FrbRGB oracle_get_recursive_pixel(const FrbConfig *cfg, uint32_t x, uint32_t y)
{
    return oracle_get_recursive_pixel_raw(cfg, x, y, NULL, NULL);
}

/*
 * get_image, Mandelbrot|Julia arm — src/lib.rs:255-269: rows in parallel (rayon work stealing at row
 * granularity -> OpenMP schedule(dynamic,1)), x serial inside a row, rows concatenated in y order.
 *
 * Generalised to an explicit row list so that tests can sample huge frames and shards:
 *   rows[k] = image row rendered into output row k (rows == NULL: row k = k, nrows = height).
 * Any of rgb (3 B/px), iters (u32/px), zfinal (2 f64/px) may be NULL. threads <= 0: all cores.
 * Returns the total number of loop-body executions (i+1 for a pixel that escaped at index i, else iterations).
 */
uint64_t oracle_render_rows(const FrbConfig *cfg, const uint32_t *rows, uint32_t nrows, uint8_t *rgb,
                            uint32_t *iters, double *zfinal, int threads)
{
    const uint32_t width = cfg->width;
    uint64_t total = 0;
    const int is_escape_algo = (cfg->algo == FRB_ALGO_MANDELBROT || cfg->algo == FRB_ALGO_JULIA);
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads) reduction(+ : total)
#endif
    for (int64_t k = 0; k < (int64_t)nrows; k++) {
        const uint32_t y = rows ? rows[k] : (uint32_t)k;
        uint64_t row_total = 0;
        for (uint32_t x = 0; x < width; x++) {
            uint32_t n;
            double z[2];
            FrbRGB c = oracle_get_recursive_pixel_raw(cfg, x, y, &n, z);
            const size_t p = (size_t)k * width + x;
            if (rgb) { rgb[3 * p] = c.r; rgb[3 * p + 1] = c.g; rgb[3 * p + 2] = c.b; }
            if (iters) iters[p] = n;
            if (zfinal) { zfinal[2 * p] = z[0]; zfinal[2 * p + 1] = z[1]; }
            if (is_escape_algo) row_total += (n < cfg->iterations) ? (uint64_t)n + 1u : (uint64_t)cfg->iterations;
        }
        total += row_total;
    }
    (void)threads;
    return total;
}

/* get_image(&Config) -> Vec<RGB> — src/lib.rs:253. */
uint64_t oracle_get_image(const FrbConfig *cfg, uint8_t *rgb, int threads)
{
    return oracle_render_rows(cfg, NULL, cfg->height, rgb, NULL, NULL, threads);
}

/* Config::new(algo) — calc/src/lib.rs:39-69 (independent of the product's frb_config_default). */
void oracle_config_new(FrbConfig *cfg, uint32_t algo)
{
    memset(cfg, 0, sizeof *cfg);
    const int fern = (algo == FRB_ALGO_BARNSLEY_FERN);
    cfg->algo = algo;
    cfg->width = 2000;
    cfg->height = 1000;
    cfg->iterations = fern ? 10000000u : 50u;
    cfg->limit = 65536.0; /* 2.0_f64.powi(16) */
    cfg->stable_limit = 2.0;
    cfg->pos.re = 0.0;
    cfg->pos.im = 0.0;
    cfg->scale.re = 1.0 * 0.4; /* Imaginary::ONE * 0.4 */
    cfg->scale.im = 1.0 * 0.4;
    cfg->exposure = 2.0;
    cfg->inside = 1;
    cfg->smooth = 1;
    cfg->primary_color = fern ? rgb_new(4, 100, 3) : rgb_new(40, 40, 255);
    cfg->secondary_color = fern ? rgb_new(240, 240, 240) : rgb_new(240, 170, 0);
    cfg->color_weight = 0.01;
    cfg->julia_set.re = 0.0;
    cfg->julia_set.im = 0.0;
}

int oracle_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
