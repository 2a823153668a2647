// fractal_renderer.hpp — C++ host-side mirror of the reference's Rust API for the escape-time path, above the C ABI.
//
// The reference is compiled code (Rust) and its toolchain is absent on this image, so the host layer a Rust caller
// would use is restated in C++ with the same names and meaning:
//   calc::Config / Imaginary / RGB / Algo   (calc/src/lib.rs:19-179)  -> fractal_renderer::calc::*
//   get_image(&Config) -> Vec<RGB>          (src/lib.rs:253)           -> fractal_renderer::get_image(const Config&)
// Errors: the reference panics (expect/unwrap); here a failed render throws std::runtime_error with frb_last_error().
// Header-only; link with libfrb.so.  No pixel is ever computed on the host.
#ifndef FRACTAL_RENDERER_HPP_
#define FRACTAL_RENDERER_HPP_

#include <cstdint>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>

#include "frb.h"

namespace fractal_renderer {
namespace calc {

enum class Algo : uint32_t { Mandelbrot = FRB_ALGO_MANDELBROT, BarnsleyFern = FRB_ALGO_BARNSLEY_FERN, Julia = FRB_ALGO_JULIA };

struct Imaginary {
    double re = 0.0, im = 0.0;
    static constexpr Imaginary ZERO() { return Imaginary{0.0, 0.0}; }
    Imaginary operator*(double rhs) const { return Imaginary{re * rhs, im * rhs}; }
};

// fields as STORED (r, g, b); RGB::new_(r, b, g) keeps the reference constructor's argument order (calc/src/lib.rs:129-131)
struct RGB {
    uint8_t r = 0, g = 0, b = 0;
    static constexpr RGB new_(uint8_t r, uint8_t b, uint8_t g) { return RGB{r, g, b}; }
    static constexpr RGB BLACK() { return RGB{0, 0, 0}; }
};
static_assert(sizeof(RGB) == 3, "Vec<RGB> is reinterpreted as bytes (src/lib.rs:13-15, src/gui.rs:62-64)");

// calc::Config with Config::new's defaults (calc/src/lib.rs:39-69)
struct Config {
    Algo algo = Algo::Mandelbrot;
    uint32_t width = 2000, height = 1000, iterations = 50;
    double limit = 65536.0, stable_limit = 2.0;
    Imaginary pos{0.0, 0.0}, scale{0.4, 0.4};
    double exposure = 2.0;
    bool inside = true, smooth = true;
    RGB primary_color = RGB::new_(40, 40, 255), secondary_color = RGB::new_(240, 170, 0);
    double color_weight = 0.01;
    Imaginary julia_set{0.0, 0.0};

    static Config new_(Algo algo)
    {
        FrbConfig c;
        frb_config_default(&c, static_cast<uint32_t>(algo));
        return from_c(c);
    }
    FrbConfig to_c() const
    {
        FrbConfig c{};
        c.algo = static_cast<uint32_t>(algo);
        c.width = width;
        c.height = height;
        c.iterations = iterations;
        c.limit = limit;
        c.stable_limit = stable_limit;
        c.pos = {pos.re, pos.im};
        c.scale = {scale.re, scale.im};
        c.exposure = exposure;
        c.inside = inside ? 1 : 0;
        c.smooth = smooth ? 1 : 0;
        c.primary_color = {primary_color.r, primary_color.g, primary_color.b};
        c.secondary_color = {secondary_color.r, secondary_color.g, secondary_color.b};
        c.color_weight = color_weight;
        c.julia_set = {julia_set.re, julia_set.im};
        return c;
    }
    static Config from_c(const FrbConfig &c)
    {
        Config k;
        k.algo = static_cast<Algo>(c.algo);
        k.width = c.width;
        k.height = c.height;
        k.iterations = c.iterations;
        k.limit = c.limit;
        k.stable_limit = c.stable_limit;
        k.pos = {c.pos.re, c.pos.im};
        k.scale = {c.scale.re, c.scale.im};
        k.exposure = c.exposure;
        k.inside = c.inside != 0;
        k.smooth = c.smooth != 0;
        k.primary_color = {c.primary_color.r, c.primary_color.g, c.primary_color.b};
        k.secondary_color = {c.secondary_color.r, c.secondary_color.g, c.secondary_color.b};
        k.color_weight = c.color_weight;
        k.julia_set = {c.julia_set.re, c.julia_set.im};
        return k;
    }
};

}  // namespace calc

inline void check(int rc, const char *what)
{
    if (rc != FRB_OK) throw std::runtime_error(std::string(what) + ": " + frb_last_error());
}

// get_image's BarnsleyFern arm (src/lib.rs:271-319): `threads` = the reference's rayon thread count (0 = this host's),
// `seed` replaces SmallRng::from_entropy (src/lib.rs:428).
inline std::vector<calc::RGB> get_image_fern(const calc::Config &config, uint32_t threads, uint64_t seed, int device = 0)
{
    std::vector<calc::RGB> image(static_cast<size_t>(config.width) * config.height);
    const FrbConfig c = config.to_c();
    check(frb_render_fern(&c, threads, seed, nullptr, reinterpret_cast<uint8_t *>(image.data()), image.size() * 3, FRB_RGB24,
                          device < 0 ? 0 : device), "get_image (fern)");
    return image;
}

// get_image(&Config) -> Vec<RGB> (src/lib.rs:253), dispatching on the algorithm as the reference does: Mandelbrot | Julia ->
// the per-pixel path (device = -1 spreads the frame over all GPUs), BarnsleyFern -> the chaos game with fresh entropy.
inline std::vector<calc::RGB> get_image(const calc::Config &config, int device = 0)
{
    if (config.algo == calc::Algo::BarnsleyFern) {
        std::random_device entropy;
        const uint64_t seed = (static_cast<uint64_t>(entropy()) << 32) ^ entropy();
        return get_image_fern(config, 0, seed, device);
    }
    std::vector<calc::RGB> image(static_cast<size_t>(config.width) * config.height);
    const FrbConfig c = config.to_c();
    check(frb_render(&c, reinterpret_cast<uint8_t *>(image.data()), image.size() * 3, FRB_RGB24, device), "get_image");
    return image;
}

// The RGBA8 frame the GUI builds from get_image (src/gui.rs:62-75), produced directly on the device.
inline std::vector<uint8_t> get_image_rgba(const calc::Config &config, int device = 0)
{
    std::vector<uint8_t> image(static_cast<size_t>(config.width) * config.height * 4);
    const FrbConfig c = config.to_c();
    check(frb_render(&c, image.data(), image.size(), FRB_RGBA32, device), "get_image_rgba");
    return image;
}

// A page-locked RGBA frame for callers that render many frames into the same memory (the GUI's render thread,
// src/gui.rs:56-82): the GPU writes a pinned destination while it iterates, so nothing is copied after the frame.
class PinnedFrame {
public:
    PinnedFrame() = default;
    PinnedFrame(const PinnedFrame &) = delete;
    PinnedFrame &operator=(const PinnedFrame &) = delete;
    ~PinnedFrame() { frb_host_free(ptr_); }
    // get_image + the RGB -> RGBA expansion of src/gui.rs:62-75; device -1 = every GPU of the box
    void render(const calc::Config &config, int device = -1)
    {
        const size_t n = (size_t)config.width * config.height * 4;
        if (n > cap_) {
            frb_host_free(ptr_);
            cap_ = n + n / 8 + 4096;
            ptr_ = static_cast<uint8_t *>(frb_host_alloc(cap_));
            if (!ptr_) {
                cap_ = 0;
                check(FRB_ERR_CUDA, "frb_host_alloc");
            }
        }
        const FrbConfig c = config.to_c();
        check(frb_render(&c, ptr_, n, FRB_RGBA32, device), "frb_render");
        len_ = n;
    }
    const uint8_t *data() const { return ptr_; }
    size_t size() const { return len_; }

private:
    uint8_t *ptr_ = nullptr;
    size_t cap_ = 0, len_ = 0;
};

}  // namespace fractal_renderer
#endif
