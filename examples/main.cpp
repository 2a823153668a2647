// main.cpp — the same consumer through the C++ host mirror (include/fractal_renderer.hpp): calc::Config, get_image(&Config)
// -> std::vector<RGB> with the reference's names (src/lib.rs:253, src/main.rs:16), errors as exceptions where Rust panics.
//
//   usage: frb_main_cpp <width> <height> <iterations> [mandelbrot|julia]      (raw RGB24 on stdout)
//   build: g++ -std=c++17 -O2 -I include examples/main.cpp -L fractal_renderer_b200 -lfrb -Wl,-rpath,'$ORIGIN/../fractal_renderer_b200' -o examples/frb_main_cpp
// Exit status: 0 = rendered; 3 = no CUDA device (there is no CPU path); 1 = any other failure.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>

#include "fractal_renderer.hpp"

int main(int argc, char **argv)
{
    using namespace fractal_renderer;
    const bool julia = argc > 4 && std::strcmp(argv[4], "julia") == 0;
    calc::Config config = calc::Config::new_(julia ? calc::Algo::Julia : calc::Algo::Mandelbrot);
    config.width = argc > 1 ? (uint32_t)std::strtoul(argv[1], nullptr, 10) : 750u;
    config.height = argc > 2 ? (uint32_t)std::strtoul(argv[2], nullptr, 10) : 500u;
    if (argc > 3) config.iterations = (uint32_t)std::strtoul(argv[3], nullptr, 10);
    config.exposure = 5.0;
    config.pos.re = julia ? 0.0 : -0.6;
    if (julia) config.julia_set = {-0.7269, 0.1889};
    static_assert(sizeof(calc::RGB) == 3, "Vec<RGB> as bytes");
    try {
        const std::vector<calc::RGB> image = get_image(config, -1);
        std::fwrite(image.data(), 3, image.size(), stdout);
        // the GUI's RGBA frame (src/gui.rs:62-75) must be the same pixels with alpha 255
        const std::vector<uint8_t> rgba = get_image_rgba(config, 0);
        for (size_t i = 0; i < image.size(); ++i)
            if (rgba[4 * i] != image[i].r || rgba[4 * i + 1] != image[i].g || rgba[4 * i + 2] != image[i].b || rgba[4 * i + 3] != 255) {
                std::fprintf(stderr, "RGBA frame differs from the RGB image at pixel %zu\n", i);
                return 1;
            }
    } catch (const std::exception &e) {
        std::fprintf(stderr, "%s\n", e.what());
        return frb_device_count() == 0 ? 3 : 1;
    }
    frb_shutdown();
    return 0;
}
