#!/bin/bash
# Writes the SASS of one instantiation of every kernel family the library ships to profiles/, with a count of
# the mnemonics that matter per kernel (UTMALDG / SYNCS = TMA + mbarrier, IDP.4A = dp4a, LDG/STG widths, SHFL, VOTE).
set -e
SO=docker-nvidia-glx-desktop_b200/libb200vc.so
OUT=profiles/sass_all_kernels.txt
: > $OUT
cuobjdump -sass $SO | sed -E 's/\s*\/\* 0x[0-9a-f]+ \*\///' | grep -vE '^\s*$' > /tmp/all.sass
for pat in \
  'k1_convert_kernelILi0ELi1ELb0ELb0ELi8ELi1ELb0E|K1 <I420,mpeg2>: bench workload C1' \
  'k1_convert_kernelILi1ELi1ELb1ELb0ELi8ELi1ELb0E|K1 <NV12,mpeg2,negated matrix>: config C2' \
  'k1_convert_kernelILi0ELi1ELb0ELb1ELi8ELi1ELb0E|K1+K3 <I420,mpeg2,damage>: config C4' \
  'k1_convert_kernelILi0ELi1ELb0ELb0ELi8ELi2ELb0E|K1 row-gated (B200VC_FLAG_ROW_GATE)' \
  'k1_convert_kernelILi3ELi0ELb0ELb0ELi8ELi1ELb1E|K1 <Y444, 5-6-5 input>' \
  'k1_convert_kernelILi2ELi1ELb0ELb0ELi8ELi1ELb0E|K1 <Y42B,mpeg2>' \
  'k2_fused_2x_kernelILi0ELi1ELb0ELb0E|K2 exact 2:1 fused <I420,mpeg2>: config C3' \
  'k2_fused_scale_kernelILi0ELi1ELb0ELi48ELb0ELb0E|K2 general fused <I420,mpeg2,48-row tiles>' \
  'k2_fused_scale_kernelILi0ELi1ELb0ELi8ELb1ELb0E|K2 general fused <I420,mpeg2,8-row tiles,odd frames>' \
  'k2_fused_scale_kernelILi0ELi1ELb0ELi32ELb0ELb1E|K2 general fused <I420,mpeg2,32-row tiles, 5-6-5 input>' \
  'k2_plane_tile_kernelILi0E|two-pass, pass 2, tiled plane scaler <planar>' \
  'k2_plane_tile_kernelILi1E|two-pass, pass 2, tiled plane scaler <NV12>' \
  'k2_plane_scale_kernelILi0E|two-pass, pass 2, gather plane scaler <planar>' \
  'static_plane_scale_kernelILi0E|no-allocation entry point, pass 2' \
  'border_fill_kernel|add-borders fill'; do
  sym=${pat%%|*}; title=${pat##*|}
  echo "==== $title ($sym)" >> $OUT
  awk -v pat="$sym" '$0 ~ "Function : .*" pat {p=1} p && /Function :/ && $0 !~ pat {p=0} p' /tmp/all.sass > /tmp/one.sass
  cat /tmp/one.sass >> $OUT
  echo "---- mnemonics of $sym" >> $OUT
  grep -oE '(UTMALDG[.A-Z0-9]*|UTMASTG[.A-Z0-9]*|UBLKCP[.A-Z0-9]*|SYNCS[.A-Z0-9]*|IDP\.4A[.A-Z0-9]*|LDG\.E[.A-Z0-9]*|STG\.E[.A-Z0-9]*|LDS[.A-Z0-9]*|STS[.A-Z0-9]*|SHFL[.A-Z]*|VOTE[.A-Z]*|PRMT|BAR\.SYNC[.A-Z_]*)' /tmp/one.sass | sort | uniq -c >> $OUT
done
wc -l $OUT
grep -n "^==== \|UTMALDG\|SYNCS.ARRIVE" $OUT | grep -v "^\S*:\s*/\*" | head -60
