/*
 * synth_image.h — deterministic synthetic test picture (SURVEY.md §8(d), config 2 recipe).
 * 5 octaves of bilinear value noise, cell sizes {256,64,16,4,1}, amplitudes {96,48,24,12,6},
 * integer-hash lattice, no libc RNG.  Output: interleaved RGB, 8 bit.
 */
#ifndef SYNTH_IMAGE_H_
#define SYNTH_IMAGE_H_
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

static inline uint32_t synth_hash(uint32_t x)
{
    x ^= x >> 16;
    x *= 0x7feb352du;
    x ^= x >> 15;
    x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}

/* hc = hash(hash(seed + octave * 0x9e3779b9) + channel): the part that does not depend on the cell */
static inline float synth_lattice_hc(uint32_t hc, int cx, int cy)
{
    uint32_t h = synth_hash(hc + (uint32_t)cx);
    h = synth_hash(h + (uint32_t)cy);
    return (float)(h & 0xffff) / 65535.0f - 0.5f;
}

static inline float synth_lattice(uint32_t seed, int octave, int channel, int cx, int cy)
{
    uint32_t h = synth_hash(seed + (uint32_t)octave * 0x9e3779b9u);
    h = synth_hash(h + (uint32_t)channel);
    return synth_lattice_hc(h, cx, cy);
}

/* amp_scale multiplies all amplitudes (1.0 = the recipe). Returns malloc'ed w*h*3 bytes. */
static uint8_t* synth_image_rgb(int w, int h, uint32_t seed, float amp_scale)
{
    static const int cells[5] = {256, 64, 16, 4, 1};
    static const float amps[5] = {96.f, 48.f, 24.f, 12.f, 6.f};
    uint8_t* img = (uint8_t*)malloc((size_t)w * h * 3);
    float* row = (float*)malloc(sizeof(float) * (size_t)w * 3);
    if (!img || !row) return 0;
    for (int y = 0; y < h; ++y) {
        for (int i = 0; i < w * 3; ++i) row[i] = 128.0f;
        for (int o = 0; o < 5; ++o) {
            int cell = cells[o];
            int cy = y / cell;
            float fy = (float)(y % cell) / (float)cell;
            for (int ch = 0; ch < 3; ++ch) {
                const uint32_t hc = synth_hash(synth_hash(seed + (uint32_t)o * 0x9e3779b9u) + (uint32_t)ch);
                if (cell == 1) {
                    /* both interpolation weights are zero: the sum below reduces to the lattice
                     * value itself, bit for bit (x + (y - x) * 0.0f == x for finite x, y) */
                    for (int x = 0; x < w; ++x)
                        row[x * 3 + ch] += amps[o] * amp_scale * synth_lattice_hc(hc, x, cy);
                    continue;
                }
                int last_cx = -1;
                float v00 = 0, v10 = 0, v01 = 0, v11 = 0;
                for (int x = 0; x < w; ++x) {
                    int cx = x / cell;
                    if (cx != last_cx) {
                        v00 = synth_lattice_hc(hc, cx, cy);
                        v10 = synth_lattice_hc(hc, cx + 1, cy);
                        v01 = synth_lattice_hc(hc, cx, cy + 1);
                        v11 = synth_lattice_hc(hc, cx + 1, cy + 1);
                        last_cx = cx;
                    }
                    float fx = (float)(x % cell) / (float)cell;
                    float top = v00 + (v10 - v00) * fx, bot = v01 + (v11 - v01) * fx;
                    row[x * 3 + ch] += amps[o] * amp_scale * (top + (bot - top) * fy);
                }
            }
        }
        for (int i = 0; i < w * 3; ++i) {
            float v = floorf(row[i] + 0.5f);
            img[(size_t)y * w * 3 + i] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
        }
    }
    free(row);
    return img;
}
#endif
