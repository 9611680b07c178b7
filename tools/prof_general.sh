bash tools/gpu_prof.sh k2_fused_scale "1440p->1080p 4:3" g43v4
bash tools/gpu_prof.sh k2_fused_scale "2:3 upscale" gupv4
