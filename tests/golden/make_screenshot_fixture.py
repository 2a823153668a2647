"""Decodes the reference's only artefact for this path — screenshots/mandelbrot-1000000x.avif — into a lossless
PNG fixture.  Run HERE (the container with /root/reference); the PNG is committed because /root/reference does not
exist on the GPU box.  The AVIF is lossy (YCbCr, quality 100), so the oracle is compared with it at
interior-mask IoU >= 0.9999 and mean |delta| <= 1 LSB per channel, not bit for bit (tests/test_oracle.py)."""
import os
import sys

from PIL import Image

SRC = "/root/reference/screenshots/mandelbrot-1000000x.avif"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "screenshot_mandelbrot_1000000x.png")

if __name__ == "__main__":
    if not os.path.exists(SRC):
        sys.exit(f"{SRC} not found: run this in the build container")
    Image.open(SRC).convert("RGB").save(DST, "PNG", optimize=True)
    print("wrote", DST, os.path.getsize(DST), "bytes")
