#!/bin/bash
# Writes the SASS of the kernels the BASELINE configs run (C1, C2, C3, C4) to profiles/.
set -e
SO=docker-nvidia-glx-desktop_b200/libb200vc.so
OUT=profiles/sass_config_kernels.txt
: > $OUT
for pat in 'k1_convert_kernelILi0ELi1ELb0ELb0ELi8' 'k1_convert_kernelILi1ELi1ELb1ELb0ELi8' 'k1_convert_kernelILi0ELi1ELb0ELb1ELi8' 'k2_fused_2x_kernelILi0ELi1ELb0ELb0'; do
  cuobjdump -sass $SO | awk -v pat="$pat" '$0 ~ "Function : .*" pat {p=1} p && /Function :/ && $0 !~ pat {p=0} p' | sed -E 's/\s*\/\* 0x[0-9a-f]+ \*\///' | grep -vE '^\s*$' >> $OUT
done
echo "== mnemonic evidence ==" >> $OUT
grep -oE '(UTMALDG[.A-Z0-9]*|SYNCS[.A-Z0-9]*|IDP\.4A[.A-Z0-9]*|LDG\.E[.A-Z0-9]*|STG\.E[.A-Z0-9]*|SHFL[.A-Z]*|VOTE[.A-Z]*|PRMT)' $OUT | sort | uniq -c >> $OUT
wc -l $OUT
