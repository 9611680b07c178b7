#!/bin/bash
# GPU box: ncu --set full captures of the general fused kernel at three ratios (tools/gpu_prof.sh).
bash tools/gpu_prof.sh k2_fused_scale "1440p->1080p 4:3" g43
bash tools/gpu_prof.sh k2_fused_scale "4K->1440p 3:2" g32
bash tools/gpu_prof.sh k2_fused_scale "2:3 upscale" gup
