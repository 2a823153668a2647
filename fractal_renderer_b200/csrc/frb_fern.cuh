// frb_fern.cuh — the Barnsley-fern arm of the reference's get_image (src/lib.rs:271-319, 366-463) on the GPU.
// SURVEY §8(f) rank 4.  Included by frb_api.cu.
//
// The reference: T = rayon threads; each thread owns an image initialised to secondary_color, plays the chaos game for
// iterations / T points from (pos.re * width, pos.im * height) with an entropy-seeded xoshiro256++ (rand 0.8 SmallRng),
// darkens the pixel under every point (subtract_pixel: multiplicative, truncating, with RGB::new's g/b swap) and the T
// images are reduced with a saturating add.  Two facts make this a GPU job without changing what is computed:
//   * a pixel's colour in one thread's image depends only on HOW MANY points hit it (every hit applies the same map), so
//     a thread's image is a hit-count histogram pushed through a table  n -> map^n(secondary_color);
//   * the reference itself splits the orbit into T independent ones ("this fractal is random, so we should
//     probabilistically get the same result", src/lib.rs:291-292); we split each of those T shares further into C chains.
//
//   F0 frb_fern_lut_kernel     table[n], n < 1024: the per-hit map iterated n times (period <= 2 from n = 510 on, see below)
//   F1 frb_fern_chaos_kernel   one thread per chain: own xoshiro256++ stream, points -> atomicAdd on counts[t][pixel]
//   F2 frb_fern_compose_kernel per pixel: saturating sum over t of table[counts[t][pixel]] -> RGB24 / RGBA32
//
// Determinism: chain j (global index) is seeded with seed + (j+1) * 0xD1342543DE82EF95 through rand_core's
// seed_from_u64 (splitmix64); chains after the first of a share discard 64 burn-in points (the first chain plots the
// transient from the start point exactly as a reference thread does).  Integer hit counts do not depend on the order of
// the atomics, so the image is a pure function of (Config, T, seed) — oracle/fern_oracle.c restates it bit for bit.
// All FP64 arithmetic uses the *_rn intrinsics in the reference's operation order (no FMA contraction).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace frb {

constexpr uint32_t kFernBurnIn = 64;
constexpr uint32_t kFernLut = 1024;
constexpr int kFernThreads = 128;

struct FernParams {
    uint32_t width, height;
    uint32_t threads;       // T: virtual reference threads
    uint32_t per;           // iterations / T (src/lib.rs:286)
    uint32_t chains;        // C: chains per virtual thread
    uint32_t format;        // 3 = RGB24, 4 = RGBA32
    uint32_t rows, band_rows, shard_index, shard_count;  // rows of the image this launch composes (FrbShard)
    uint32_t primary, secondary;  // stored fields r | g<<8 | b<<16
    uint64_t seed;
    uint64_t npix;
    double pos_re, pos_im;
    double wf, hf;          // width, height as f64
    double esx, esy;        // effective_scale_x / _y (src/lib.rs:425-426)
    double amount;          // color_weight
    uint32_t *counts;       // [T][npix]
    uint32_t *lut;          // [kFernLut] packed r | g<<8 | b<<16
    uint8_t *out;
};

// Rust `f64 as usize`: truncate toward zero, saturate to [0, 2^64-1], NaN -> 0.
__device__ __forceinline__ uint64_t fern_sat_usize(double v)
{
    if (!(v == v)) return 0ull;
    if (v <= 0.0) return 0ull;
    if (v >= 18446744073709551616.0) return 0xffffffffffffffffull;
    return __double2ull_rz(v);
}

__device__ __forceinline__ uint32_t fern_sat_u8(double v)
{
    return (v == v) ? min(__double2uint_rz(v), 255u) : 0u;
}

// subtract_pixel (src/lib.rs:388-408) for one channel: pixel as f64 * 1.0 / ((((1.0 / (value / 255.0)) - 1.0) * amount) + 1.0)
__device__ __forceinline__ uint32_t fern_channel(uint32_t pixel, uint32_t value, double amount)
{
    const double d = __dadd_rn(__dmul_rn(__dsub_rn(__ddiv_rn(1.0, __ddiv_rn((double)value, 255.0)), 1.0), amount), 1.0);
    return fern_sat_u8(__ddiv_rn(__dmul_rn((double)pixel, 1.0), d));
}

// F0: table[n] = colour of a pixel of one thread's image after n hits.  With D_c - 1 = (255/value_c - 1) * amount all three
// channel maps move in the same direction (sign of amount), so r and the two-hit maps of g and b are monotone on 0..255
// and the state has period <= 2 after at most 510 hits: for n >= 1024 the compose kernel reads table[1022 + (n & 1)].
__global__ void frb_fern_lut_kernel(const __grid_constant__ FernParams p)
{
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    uint32_t r = p.secondary & 0xffu, g = (p.secondary >> 8) & 0xffu, b = (p.secondary >> 16) & 0xffu;
    const uint32_t vr = p.primary & 0xffu, vg = (p.primary >> 8) & 0xffu, vb = (p.primary >> 16) & 0xffu;
    for (uint32_t n = 0; n < kFernLut; ++n) {
        p.lut[n] = r | (g << 8) | (b << 16);
        const uint32_t nr = fern_channel(r, vr, p.amount);
        const uint32_t ng = fern_channel(g, vg, p.amount);
        const uint32_t nb = fern_channel(b, vb, p.amount);
        // RGB::new(nr, ng, nb) with new(r, b, g) (calc/src/lib.rs:129-131): the second argument lands in .b, the third in .g
        r = nr;
        b = ng;
        g = nb;
    }
}

// xoshiro256++ (rand 0.8 SmallRng on 64-bit targets), seeded like rand_core's seed_from_u64 (splitmix64).
struct Xoshiro {
    uint64_t s0, s1, s2, s3;
};

__device__ __forceinline__ uint64_t rotl64(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }

__device__ __forceinline__ void xoshiro_seed(Xoshiro &g, uint64_t state)
{
    uint64_t out[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        state += 0x9E3779B97F4A7C15ull;
        uint64_t z = state;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        out[i] = z ^ (z >> 31);
    }
    g.s0 = out[0];
    g.s1 = out[1];
    g.s2 = out[2];
    g.s3 = out[3];
}

// `rng.gen::<f64>()`: 53 bits of next_u64 scaled into [0, 1)
__device__ __forceinline__ double xoshiro_f64(Xoshiro &g)
{
    const uint64_t result = rotl64(g.s0 + g.s3, 23) + g.s0;
    const uint64_t t = g.s1 << 17;
    g.s2 ^= g.s0;
    g.s3 ^= g.s1;
    g.s1 ^= g.s2;
    g.s0 ^= g.s3;
    g.s2 ^= t;
    g.s3 = rotl64(g.s3, 45);
    return __dmul_rn((double)(result >> 11), 1.0 / 9007199254740992.0);
}

// F1: the chaos game (src/lib.rs:417-463).  Thread = chain; chain j serves virtual thread j / C.
__global__ void __launch_bounds__(kFernThreads)
frb_fern_chaos_kernel(const __grid_constant__ FernParams p)
{
    // the four affine maps, x' = A x + B y, y' = (C old_x + D y) + E, coefficients exactly as written in the reference
    // (0.20 * x - 0.26 * y == 0.20 * x + (-0.26) * y in IEEE arithmetic)
    __shared__ double coef[4][5];
    if (threadIdx.x == 0) {
        const double tab[4][5] = {{0.00, 0.00, 0.00, 0.16, 0.00},
                                  {0.85, 0.04, -0.04, 0.85, 1.60},
                                  {0.20, -0.26, 0.23, 0.22, 1.60},
                                  {-0.15, 0.28, 0.26, 0.24, 0.44}};
        for (int i = 0; i < 4; ++i)
            for (int j = 0; j < 5; ++j) coef[i][j] = tab[i][j];
    }
    __syncthreads();

    const uint64_t chain = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (chain >= (uint64_t)p.threads * p.chains) return;
    const uint32_t t = (uint32_t)(chain / p.chains);
    const uint32_t k = (uint32_t)(chain - (uint64_t)t * p.chains);
    const uint32_t steps = p.per / p.chains + (k < p.per % p.chains ? 1u : 0u);

    Xoshiro g;
    xoshiro_seed(g, p.seed + (chain + 1ull) * 0xD1342543DE82EF95ull);
    double x = __dmul_rn(p.pos_re, p.wf), y = __dmul_rn(p.pos_im, p.hf);  // src/lib.rs:421-422
    const double half_w = __ddiv_rn(p.wf, 2.0), half_h = __ddiv_rn(p.hf, 2.0);
    const double pim5 = __dsub_rn(p.pos_im, 5.0);
    uint32_t *const counts = p.counts + (size_t)t * p.npix;

    const uint32_t burn = k > 0 ? kFernBurnIn : 0u;
    for (uint32_t i = 0; i < burn + steps; ++i) {
        if (i >= burn) {
            // subtract_pixel's coordinates (src/lib.rs:433-438) and pixel_mut's bounds (src/lib.rs:378-387)
            const uint64_t px = fern_sat_usize(__dadd_rn(__dmul_rn(__dsub_rn(x, p.pos_re), p.esx), half_w));
            const double v = __dadd_rn(__dmul_rn(__dsub_rn(__dadd_rn(y, pim5), 0.5), p.esy), half_h);
            const uint64_t py = fern_sat_usize(__dsub_rn(p.hf, v));
            if (px <= (uint64_t)p.width) {  // `x > self.width` rejects; x == width lands on the next row's first pixel
                const uint64_t index = py * (uint64_t)p.width + px;  // wrapping usize arithmetic
                if (index < p.npix) atomicAdd(counts + index, 1u);
            }
        }
        const double r = xoshiro_f64(g);
        const int m = (r < 0.01 ? 0 : 1) + (r < 0.86 ? 0 : 1) + (r < 0.93 ? 0 : 1);
        const double old_x = x;
        x = __dadd_rn(__dmul_rn(coef[m][0], x), __dmul_rn(coef[m][1], y));
        y = __dadd_rn(__dadd_rn(__dmul_rn(coef[m][2], old_x), __dmul_rn(coef[m][3], y)), coef[m][4]);
    }
}

// F2: saturating reduction of the T images (src/lib.rs:274-284, 301-318; calc/src/lib.rs:140-146), one pixel per thread.
__global__ void __launch_bounds__(256)
frb_fern_compose_kernel(const __grid_constant__ FernParams p)
{
    const uint64_t n_out = (uint64_t)p.rows * p.width;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_out; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t lr = (uint32_t)(i / p.width);
        const uint32_t px = (uint32_t)(i - (uint64_t)lr * p.width);
        const uint32_t band = lr / p.band_rows;
        const uint32_t y = (band * p.shard_count + p.shard_index) * p.band_rows + (lr - band * p.band_rows);
        const uint64_t pix = (uint64_t)y * p.width + px;
        uint32_t r = 0, g = 0, b = 0;
        for (uint32_t t = 0; t < p.threads; ++t) {
            const uint32_t n = p.counts[(size_t)t * p.npix + pix];
            const uint32_t c = p.lut[n < kFernLut ? n : (kFernLut - 2u) + (n & 1u)];
            r = min(r + (c & 0xffu), 255u);
            g = min(g + ((c >> 8) & 0xffu), 255u);
            b = min(b + ((c >> 16) & 0xffu), 255u);
        }
        if (p.format == 4u) {
            reinterpret_cast<uint32_t *>(p.out)[i] = r | (g << 8) | (b << 16) | 0xff000000u;
        } else {
            p.out[i * 3 + 0] = (uint8_t)r;
            p.out[i * 3 + 1] = (uint8_t)g;
            p.out[i * 3 + 2] = (uint8_t)b;
        }
    }
}

}  // namespace frb
