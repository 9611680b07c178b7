#include "oracle.h"
#include <stdio.h>
#include <stdlib.h>
int main(void) {
    int sizes[][2] = {{2,2},{3,3},{5,2},{2,5},{6,4},{7,9},{130,18},{131,35},{258,34},{1366,768},{1365,767}};
    for (unsigned k = 0; k < sizeof sizes / sizeof sizes[0]; k++) {
        int w = sizes[k][0], h = sizes[k][1];
        uint8_t *in = malloc((size_t)w * h * 4), *prev = malloc((size_t)w * h * 4);
        orc_synth_desk(in, (size_t)w * 4, w, h, k); orc_synth_rand(prev, (size_t)w * 4, w, h, k);
        for (int fmt = 0; fmt < 2; fmt++) for (int site = 0; site < 2; site++) for (int inf = 0; inf < 4; inf++) {
            size_t off[3], pit[3]; size_t total = orc_default_layout(fmt, w, h, off, pit);
            uint8_t *out = malloc(total); uint8_t *pl[3] = { out + off[0], out + off[1], out + off[2] };
            if (orc_convert_scale_mt_ex(in, (size_t)w * 4, inf, w, h, w, h, fmt, 1, site, site, 0, pl, pit, 3)) return 1;
            free(out);
            int ow = w > 4 ? (w * 2 / 3) : w, oh = h > 4 ? (h * 2 / 3) : h; if (ow < 2) ow = 2; if (oh < 2) oh = 2;
            total = orc_default_layout(fmt, ow, oh, off, pit);
            out = malloc(total); uint8_t *p2[3] = { out + off[0], out + off[1], out + off[2] };
            if (orc_convert_scale_mt_ex(in, (size_t)w * 4, inf, w, h, ow, oh, fmt, 0, 1, site, 0, p2, pit, 2)) return 2;
            free(out);
        }
        uint8_t *map = malloc((size_t)((w + 15) / 16) * ((h + 15) / 16));
        orc_damage_map(in, (size_t)w * 4, prev, (size_t)w * 4, w, h, map);
        free(map); free(in); free(prev);
    }
    puts("oracle clean under ASan/UBSan");
    return 0;
}
