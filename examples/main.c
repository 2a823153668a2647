/* main.c — a plain-C stand-in for the reference's CLI binary (src/main.rs:4-23): build a Config, call what get_image binds to
 * (frb_render), write the image bytes.  Exercises the C ABI of include/frb.h from a compiled, non-Python consumer.
 *
 *   usage: frb_main <width> <height> <iterations> [mandelbrot|julia] [out.ppm | -]
 *   build: gcc -std=c11 -O2 -I include examples/main.c -L fractal_renderer_b200 -lfrb -Wl,-rpath,'$ORIGIN/../fractal_renderer_b200' -o examples/frb_main
 *
 * "-" (the default) writes the raw RGB24 bytes to stdout — what write_image would hand to the encoder (src/lib.rs:324).
 * Exit status: 0 = rendered; 3 = FRB_ERR_NO_DEVICE (there is no CPU path); 1 = any other failure. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "frb.h"

int main(int argc, char **argv)
{
    FrbConfig cfg;
    const char *algo = argc > 4 ? argv[4] : "mandelbrot";
    const char *path = argc > 5 ? argv[5] : "-";
    size_t n;
    uint8_t *image;
    int rc;

    /* the configuration the reference CLI builds with no flags (src/lib.rs:31-234): Config::new(algo) for iterations and
     * colours, 750x500, exposure 5, pos.re -0.6 (0 for julia) */
    frb_config_default(&cfg, strcmp(algo, "julia") == 0 ? FRB_ALGO_JULIA : FRB_ALGO_MANDELBROT);
    cfg.width = argc > 1 ? (uint32_t)strtoul(argv[1], NULL, 10) : 750u;
    cfg.height = argc > 2 ? (uint32_t)strtoul(argv[2], NULL, 10) : 500u;
    if (argc > 3) cfg.iterations = (uint32_t)strtoul(argv[3], NULL, 10);
    cfg.exposure = 5.0;
    cfg.pos.re = cfg.algo == FRB_ALGO_JULIA ? 0.0 : -0.6;
    if (cfg.algo == FRB_ALGO_JULIA) {
        cfg.julia_set.re = -0.7269; /* examples.md:8 */
        cfg.julia_set.im = 0.1889;
    }

    if (frb_abi_version() != FRB_ABI_VERSION) {
        fprintf(stderr, "libfrb ABI %u, header %u\n", frb_abi_version(), FRB_ABI_VERSION);
        return 1;
    }
    n = (size_t)cfg.width * cfg.height * 3;
    image = (uint8_t *)malloc(n ? n : 1); /* Vec::with_capacity: fresh, pageable */
    if (!image) return 1;
    rc = frb_render(&cfg, image, n, FRB_RGB24, -1); /* -1: every GPU of the box */
    if (rc != FRB_OK) {
        fprintf(stderr, "frb_render failed (%d): %s\n", rc, frb_last_error());
        free(image);
        return rc == FRB_ERR_NO_DEVICE ? 3 : 1;
    }
    if (strcmp(path, "-") == 0) {
        fwrite(image, 1, n, stdout);
    } else {
        FILE *f = fopen(path, "wb");
        if (!f) return 1;
        fprintf(f, "P6\n%u %u\n255\n", cfg.width, cfg.height);
        fwrite(image, 1, n, f);
        fclose(f);
    }
    {
        FrbStats st;
        if (frb_last_stats(&st) == FRB_OK)
            fprintf(stderr, "%llu pixels, %llu iterations, %.3f ms on %u device(s)\n", (unsigned long long)st.pixels,
                    (unsigned long long)st.iterations, st.total_ms, st.devices);
    }
    free(image);
    frb_shutdown();
    return 0;
}
