/*
 * damage_synth.c — 16x16 macroblock damage map and the synthetic frame generators.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).
 *
 * The damage map has no counterpart in the reference: selkies runs ximagesrc with XDamage off,
 * and the reference installs libxdamage only for x11vnc (Dockerfile:438,496, entrypoint.sh:123).
 * Its definition is SURVEY.md §8 (a7). The generators restate SURVEY §8(d) "Synthetic inputs";
 * they stand in for ximagesrc / videotestsrc, which need X and GStreamer.
 */
#include "oracle.h"

#include <string.h>

int orc_damage_map(const uint8_t *cur, size_t cur_pitch, const uint8_t *prev, size_t prev_pitch,
                   int w, int h, uint8_t *map)
{
    return orc_damage_map_ex(cur, cur_pitch, prev, prev_pitch, ORC_IN_BGRX, w, h, map);
}

int orc_damage_map_ex(const uint8_t *cur, size_t cur_pitch, const uint8_t *prev, size_t prev_pitch,
                      int in_format, int w, int h, uint8_t *map)
{
    if (!cur || !map || w < 1 || h < 1 || in_format < ORC_IN_BGRX || in_format > ORC_IN_BGR16)
        return -1;
    const int bpp16 = in_format >= ORC_IN_RGB16;          /* 2-byte pixels: every bit is colour */
    const int c0 = (in_format == ORC_IN_XRGB || in_format == ORC_IN_XBGR) ? 1 : 0;   /* first colour byte */
    const int mb_w = (w + 15) / 16, mb_h = (h + 15) / 16;
    if (!prev) {
        memset(map, 1, (size_t)mb_w * mb_h);
        return 0;
    }
    memset(map, 0, (size_t)mb_w * mb_h);
    for (int y = 0; y < h; y++) {
        const uint8_t *a = cur + (size_t)y * cur_pitch;
        const uint8_t *b = prev + (size_t)y * prev_pitch;
        for (int x = 0; x < w; x++) {
            if (bpp16) {
                if (a[2 * x] != b[2 * x] || a[2 * x + 1] != b[2 * x + 1])
                    map[(y / 16) * mb_w + (x / 16)] = 1;
                continue;
            }
            /* the x byte cannot change the converted output, so it is not damage */
            if (a[4 * x + c0] != b[4 * x + c0] || a[4 * x + c0 + 1] != b[4 * x + c0 + 1] ||
                a[4 * x + c0 + 2] != b[4 * x + c0 + 2])
                map[(y / 16) * mb_w + (x / 16)] = 1;
        }
    }
    return 0;
}

/* ---- generators -------------------------------------------------------------------------- */

uint64_t orc_splitmix64(uint64_t *state)
{
    uint64_t z = (*state += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

/* S-rand: every byte (x included) uniform; one generator per row so rows are independent of pitch. */
void orc_synth_rand(uint8_t *dst, size_t pitch, int w, int h, uint64_t frame_index)
{
    for (int y = 0; y < h; y++) {
        uint64_t st = 0x9E3779B97F4A7C15ull + frame_index * 0x100000001B3ull + (uint64_t)y * 0xD6E8FEB86659FD93ull;
        uint8_t *row = dst + (size_t)y * pitch;
        int n = w * 4, i = 0;
        for (; i + 8 <= n; i += 8) {
            uint64_t v = orc_splitmix64(&st);
            memcpy(row + i, &v, 8);
        }
        if (i < n) {
            uint64_t v = orc_splitmix64(&st);
            memcpy(row + i, &v, (size_t)(n - i));
        }
    }
}

static inline void put(uint8_t *row, int x, int r, int g, int b)
{
    row[4 * x + 0] = (uint8_t)b;
    row[4 * x + 1] = (uint8_t)g;
    row[4 * x + 2] = (uint8_t)r;
}

/* S-bars: top 2/3 eight saturated bars; next 1/6 a 0..255 ramp; rest a 1-pixel checkerboard.
 * The x byte carries a pattern so that a kernel reading it by mistake is caught. */
void orc_synth_bars(uint8_t *dst, size_t pitch, int w, int h)
{
    static const uint8_t bars[8][3] = {
        {255, 255, 255}, {255, 255, 0}, {0, 255, 255}, {0, 255, 0},
        {255, 0, 255}, {255, 0, 0}, {0, 0, 255}, {0, 0, 0},
    };
    const int y1 = (h * 2) / 3, y2 = (h * 5) / 6;
    for (int y = 0; y < h; y++) {
        uint8_t *row = dst + (size_t)y * pitch;
        for (int x = 0; x < w; x++) {
            if (y < y1) {
                int k = (int)(((long)x * 8) / w);
                put(row, x, bars[k][0], bars[k][1], bars[k][2]);
            } else if (y < y2) {
                int v = (w > 1) ? (int)(((long)x * 255) / (w - 1)) : 0;
                put(row, x, v, v, v);
            } else {
                int v = ((x + y) & 1) ? 255 : 0;
                put(row, x, v, v, v);
            }
            row[4 * x + 3] = (uint8_t)(x * 7 + y * 13);
        }
    }
}

static void fill_rect(uint8_t *dst, size_t pitch, int w, int h, int x0, int y0, int rw, int rh,
                      int r, int g, int b)
{
    for (int y = y0; y < y0 + rh; y++) {
        if (y < 0 || y >= h)
            continue;
        uint8_t *row = dst + (size_t)y * pitch;
        for (int x = x0; x < x0 + rw; x++)
            if (x >= 0 && x < w)
                put(row, x, r, g, b);
    }
}

/* S-desk: bars base + moving 16x24 cursor + 2x16 caret toggling every 30 frames + 320x180 noise. */
void orc_synth_desk(uint8_t *dst, size_t pitch, int w, int h, uint64_t frame_index)
{
    orc_synth_bars(dst, pitch, w, h);
    const int k = (int)(frame_index & 0x7fffffff);

    int cw = w > 16 ? w - 16 : 1, ch = h > 24 ? h - 24 : 1;
    int cx = (7 * k) % cw, cy = (3 * k) % ch;
    fill_rect(dst, pitch, w, h, cx, cy, 16, 24, 0, 0, 0);
    fill_rect(dst, pitch, w, h, cx + 1, cy + 1, 14, 22, 255, 255, 255);

    if ((k / 30) & 1)
        fill_rect(dst, pitch, w, h, w / 4, h / 2, 2, 16, 32, 32, 32);

    int nx = w / 2, ny = h / 4;
    uint64_t st = 0x51ED270B5F3A9C11ull ^ (frame_index * 0x9E3779B97F4A7C15ull);
    for (int y = ny; y < ny + 180 && y < h; y++) {
        uint8_t *row = dst + (size_t)y * pitch;
        for (int x = nx; x < nx + 320 && x < w; x++) {
            uint64_t v = orc_splitmix64(&st);
            put(row, x, (int)(v & 0xff), (int)((v >> 8) & 0xff), (int)((v >> 16) & 0xff));
        }
    }
}
