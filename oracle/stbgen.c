/*
 * stbgen.c — TEST INFRASTRUCTURE.  Drives the stb_image_write.h that ships with the reference
 * (util/stb_image_write.h, included from where it lies; never copied) to synthesise the input
 * files BASELINE.json names:  stbgen <out.jpg> <width> <height> <seed> [quality=90] [amp=1.0]
 * quality <= 90 gives 4:2:0, > 90 gives 4:4:4 (util/stb_image_write.h:1478-1479).
 */
#define STB_IMAGE_WRITE_IMPLEMENTATION
#include "stb_image_write.h"

#include "../tools/synth_image.h"

#include <stdio.h>

int main(int argc, char** argv)
{
    if (argc < 5) {
        fprintf(stderr, "usage: %s out.jpg width height seed [quality] [amp]\n", argv[0]);
        return 2;
    }
    int w = atoi(argv[2]), h = atoi(argv[3]);
    uint32_t seed = (uint32_t)strtoul(argv[4], 0, 10);
    int quality = argc > 5 ? atoi(argv[5]) : 90;
    float amp = argc > 6 ? (float)atof(argv[6]) : 1.0f;
    uint8_t* img = synth_image_rgb(w, h, seed, amp);
    if (!img) return 1;
    int ok = stbi_write_jpg(argv[1], w, h, 3, img, quality);
    free(img);
    return ok ? 0 : 1;
}
