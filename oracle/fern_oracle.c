/*
 * fern_oracle.c — CPU restatement of the reference's Barnsley-fern arm (TEST INFRASTRUCTURE, never the product).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may load this.  Follows, function by function:
 *   src/lib.rs:271-319   get_image, BarnsleyFern arm: T = rayon threads, per-thread iterations = iterations / T, one
 *                        image per thread initialised to secondary_color, images reduced with a saturating add
 *   src/lib.rs:366-408   Image::pixel_mut (x > width rejects, x == width wraps to the next row) and subtract_pixel
 *                        (multiplicative darkening, truncating `as u8`, RGB::new's g/b swap on every hit)
 *   src/lib.rs:417-463   fern(): the chaos game (4 affine maps, r < 0.01 / 0.86 / 0.93), point plotted BEFORE it is moved
 *   calc/src/lib.rs:129-146  RGB::new(r, b, g), AddAssign = saturating_add per channel
 *
 * PARITY UNPINNED, and not pinnable bit for bit: the reference seeds rand::rngs::SmallRng from OS entropy
 * (src/lib.rs:428) and its result depends on the host's thread count, so no two runs of the reference agree with each
 * other.  Two things are therefore restated here:
 *   (1) fern_reference_style(): the reference's exact structure — T sequential streams of iterations/T points, each a
 *       xoshiro256++ generator (what SmallRng is on 64-bit targets in rand 0.8.4, Cargo.lock) seeded from a given u64
 *       with rand_core's seed_from_u64 (splitmix64), f64 = (next_u64 >> 11) * 2^-53 (rand's Standard distribution).
 *   (2) fern_chained(): the DETERMINISTIC decomposition the CUDA path uses (include/frb.h: frb_render_fern): every
 *       virtual thread's share is cut into C chains with their own generators, chains after the first discard 64 burn-in
 *       points.  The GPU result must equal this bit for bit (integer hit counts are order independent); (2) must agree
 *       with (1) statistically (tests/test_fern.py).
 * Everything after the hit counts — the per-hit colour map iterated n times, the saturating reduction over the T images —
 * is deterministic and shared by both.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/frb.h"

#define FERN_BURN_IN 64u

/* ---- generators ------------------------------------------------------------------------------------------- */

typedef struct { uint64_t s[4]; } Xo;

static inline uint64_t rotl64(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }

/* xoshiro256++ (Blackman & Vigna), the algorithm behind rand 0.8's SmallRng on 64-bit platforms */
static inline uint64_t xo_next(Xo *g)
{
    uint64_t *s = g->s;
    const uint64_t result = rotl64(s[0] + s[3], 23) + s[0];
    const uint64_t t = s[1] << 17;
    s[2] ^= s[0];
    s[3] ^= s[1];
    s[1] ^= s[2];
    s[0] ^= s[3];
    s[2] ^= t;
    s[3] = rotl64(s[3], 45);
    return result;
}

/* rand_core 0.6 SeedableRng::seed_from_u64 for xoshiro: four splitmix64 outputs */
static void xo_seed(Xo *g, uint64_t state)
{
    for (int i = 0; i < 4; ++i) {
        state += 0x9E3779B97F4A7C15ull;
        uint64_t z = state;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        g->s[i] = z ^ (z >> 31);
    }
}

/* rand 0.8 Standard for f64: 53 random bits scaled into [0, 1) */
static inline double xo_f64(Xo *g) { return (double)(xo_next(g) >> 11) * (1.0 / 9007199254740992.0); }

/* Seed of chain `chain` (0-based, global over all virtual threads) derived from the caller's seed. */
static inline uint64_t chain_seed(uint64_t seed, uint64_t chain) { return seed + (chain + 1u) * 0xD1342543DE82EF95ull; }

/* ---- geometry --------------------------------------------------------------------------------------------- */

/* Rust `f64 as usize` on a 64-bit target: truncate toward zero, saturate, NaN -> 0. */
static inline uint64_t sat_usize(double v)
{
    if (!(v == v)) return 0;
    if (v <= 0.0) return 0;
    if (v >= 18446744073709551616.0) return UINT64_MAX;
    return (uint64_t)v;
}

typedef struct {
    double width, height, pos_re, pos_im, esx, esy;
    uint64_t w, npix;
} FernView;

static FernView fern_view(const FrbConfig *cfg)
{
    FernView v;
    v.width = (double)cfg->width;
    v.height = (double)cfg->height;
    v.pos_re = cfg->pos.re;
    v.pos_im = cfg->pos.im;
    /* src/lib.rs:424-426, left to right */
    v.esx = 65.0 * cfg->scale.re * (double)cfg->height * 0.006;
    v.esy = 37.0 * cfg->scale.im * (double)cfg->height * 0.006;
    v.w = cfg->width;
    v.npix = (uint64_t)cfg->width * cfg->height;
    return v;
}

/* The pixel a point lands on (src/lib.rs:433-438 + pixel_mut :378-387); returns npix for "no pixel". */
static inline uint64_t fern_pixel(const FernView *v, double x, double y)
{
    const uint64_t px = sat_usize(((x - v->pos_re) * v->esx) + v->width / 2.0);
    const uint64_t py = sat_usize(v->height - ((y + (v->pos_im - 5.0) - 0.5) * v->esy + v->height / 2.0));
    if (px > v->w) return v->npix;           /* `if x > self.width { return None }` — x == width passes */
    const uint64_t index = py * v->w + px;    /* usize arithmetic (wrapping in release builds) */
    if (index >= v->npix) return v->npix;     /* len < index -> None; get_mut(len) -> None */
    return index;
}

/* One step of the chaos game (src/lib.rs:441-461), written as the reference writes it. */
static inline void fern_step(double r, double *px, double *py)
{
    double x = *px, y = *py;
    const double old_x = x;
    if (r < 0.01) {
        x = 0.00 * x + 0.00 * y;
        y = 0.00 * old_x + 0.16 * y + 0.00;
    } else if (r < 0.86) {
        x = 0.85 * x + 0.04 * y;
        y = -0.04 * old_x + 0.85 * y + 1.60;
    } else if (r < 0.93) {
        x = 0.20 * x - 0.26 * y;
        y = 0.23 * old_x + 0.22 * y + 1.60;
    } else {
        x = -0.15 * x + 0.28 * y;
        y = 0.26 * old_x + 0.24 * y + 0.44;
    }
    *px = x;
    *py = y;
}

/* ---- hit counts -------------------------------------------------------------------------------------------- */

/* Chains per virtual thread for `per` points: about 512 points per chain, at least 1, at most 65536. Shared spec with the
 * CUDA path (frb_fern.cuh: fern_chains_per_thread). */
uint32_t fern_oracle_chains_per_thread(uint32_t per)
{
    uint64_t c = ((uint64_t)per + 511u) / 512u;
    if (c < 1) c = 1;
    if (c > 65536u) c = 65536u;
    return (uint32_t)c;
}

/* (2) the decomposition of the CUDA path: counts[t * npix + pixel] += 1 for each plotted point of virtual thread t. */
void fern_oracle_counts_chained(const FrbConfig *cfg, uint32_t threads, uint64_t seed, uint32_t *counts)
{
    const FernView v = fern_view(cfg);
    memset(counts, 0, (size_t)threads * v.npix * sizeof(uint32_t));
    const uint32_t per = cfg->iterations / threads; /* src/lib.rs:286 */
    const uint32_t C = fern_oracle_chains_per_thread(per);
    for (uint32_t t = 0; t < threads; ++t)
        for (uint32_t k = 0; k < C; ++k) {
            const uint32_t steps = per / C + (k < per % C ? 1u : 0u);
            Xo g;
            xo_seed(&g, chain_seed(seed, (uint64_t)t * C + k));
            double x = v.pos_re * v.width, y = v.pos_im * v.height; /* src/lib.rs:421-422 */
            if (k > 0)
                for (uint32_t i = 0; i < FERN_BURN_IN; ++i) fern_step(xo_f64(&g), &x, &y);
            for (uint32_t i = 0; i < steps; ++i) {
                const uint64_t idx = fern_pixel(&v, x, y);
                if (idx < v.npix) counts[(size_t)t * v.npix + idx]++;
                fern_step(xo_f64(&g), &x, &y);
            }
        }
}

/* (1) the reference's own structure: one generator and one uninterrupted orbit per thread. */
void fern_oracle_counts_reference_style(const FrbConfig *cfg, uint32_t threads, uint64_t seed, uint32_t *counts)
{
    const FernView v = fern_view(cfg);
    memset(counts, 0, (size_t)threads * v.npix * sizeof(uint32_t));
    const uint32_t per = cfg->iterations / threads;
    for (uint32_t t = 0; t < threads; ++t) {
        Xo g;
        xo_seed(&g, seed + t); /* stands in for SmallRng::from_entropy() */
        double x = v.pos_re * v.width, y = v.pos_im * v.height;
        for (uint32_t i = 0; i < per; ++i) {
            const uint64_t idx = fern_pixel(&v, x, y);
            if (idx < v.npix) counts[(size_t)t * v.npix + idx]++;
            fern_step(xo_f64(&g), &x, &y);
        }
    }
}

/* ---- colour -------------------------------------------------------------------------------------------------- */

static inline uint8_t sat_u8f(double v)
{
    if (!(v == v)) return 0;
    if (v <= 0.0) return 0;
    if (v >= 255.0) return 255;
    return (uint8_t)v;
}

/* subtract_pixel (src/lib.rs:388-408) applied once to `p` with value = primary_color, amount = color_weight. */
static inline FrbRGB fern_hit(FrbRGB p, FrbRGB value, double amount)
{
    const double nr = (double)p.r * 1.0 / ((((1.0 / ((double)value.r / 255.0)) - 1.0) * amount) + 1.0);
    const double ng = (double)p.g * 1.0 / ((((1.0 / ((double)value.g / 255.0)) - 1.0) * amount) + 1.0);
    const double nb = (double)p.b * 1.0 / ((((1.0 / ((double)value.b / 255.0)) - 1.0) * amount) + 1.0);
    FrbRGB out; /* RGB::new(nr as u8, ng as u8, nb as u8) with new(r, b, g): second argument lands in .b, third in .g */
    out.r = sat_u8f(nr);
    out.b = sat_u8f(ng);
    out.g = sat_u8f(nb);
    return out;
}

static inline int rgb_eq(FrbRGB a, FrbRGB b) { return a.r == b.r && a.g == b.g && a.b == b.b; }

/* The colour of a pixel of one thread's image after n hits: n applications of fern_hit to secondary_color
 * (src/lib.rs:295), cut short once the state repeats with period 1 or 2 (the g/b swap makes 2-cycles possible). */
FrbRGB fern_oracle_colour_after(const FrbConfig *cfg, uint32_t n)
{
    FrbRGB p = cfg->secondary_color, prev = p;
    for (uint32_t h = 0; h < n; ++h) {
        const FrbRGB q = fern_hit(p, cfg->primary_color, cfg->color_weight); /* state after h + 1 hits */
        if (rgb_eq(q, p)) return p;
        if (h >= 1 && rgb_eq(q, prev)) return ((n - (h + 1)) & 1u) ? p : q;
        prev = p;
        p = q;
    }
    return p;
}

/* Reduction of the T images (src/lib.rs:274-284, 301-318; calc/src/lib.rs:140-146) -> RGB24, row-major. */
void fern_oracle_compose(const FrbConfig *cfg, uint32_t threads, const uint32_t *counts, uint8_t *rgb)
{
    const uint64_t npix = (uint64_t)cfg->width * cfg->height;
    for (uint64_t i = 0; i < npix; ++i) {
        unsigned r = 0, g = 0, b = 0;
        for (uint32_t t = 0; t < threads; ++t) {
            const FrbRGB p = fern_oracle_colour_after(cfg, counts[(size_t)t * npix + i]);
            r += p.r;
            g += p.g;
            b += p.b;
        }
        rgb[3 * i + 0] = (uint8_t)(r > 255 ? 255 : r); /* saturating_add per channel; associative, so the reduce tree's shape is irrelevant */
        rgb[3 * i + 1] = (uint8_t)(g > 255 ? 255 : g);
        rgb[3 * i + 2] = (uint8_t)(b > 255 ? 255 : b);
    }
}

/* get_image, BarnsleyFern arm, with the CUDA path's decomposition (2): returns the number of plotted points. */
uint64_t fern_oracle_get_image(const FrbConfig *cfg, uint32_t threads, uint64_t seed, uint8_t *rgb, int reference_style)
{
    const uint64_t npix = (uint64_t)cfg->width * cfg->height;
    if (threads == 0 || npix == 0) return 0;
    uint32_t *counts = (uint32_t *)malloc((size_t)threads * npix * sizeof(uint32_t));
    if (!counts) return 0;
    if (reference_style) fern_oracle_counts_reference_style(cfg, threads, seed, counts);
    else fern_oracle_counts_chained(cfg, threads, seed, counts);
    fern_oracle_compose(cfg, threads, counts, rgb);
    free(counts);
    return (uint64_t)(cfg->iterations / threads) * threads;
}
