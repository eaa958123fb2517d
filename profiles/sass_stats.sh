#!/bin/bash
# SASS statistics of the solve kernels in a built library: instruction count, histogram of the mnemonics that matter for
# this FP64 kernel (DFMA/DMUL/DADD, MUFU.RCP64H, LDS/STS widths, SHFL, BAR/WARPSYNC), tensor/TMA mnemonics (none expected:
# there is no tcgen05 FP64 path), registers and spills from the ELF resource usage.
#   bash profiles/sass_stats.sh g1_locomotion_b200/libg1mpc_b200.so "ILi10ELi4E" > profiles/r2_sass_solve.txt
lib=${1:-g1_locomotion_b200/libg1mpc_b200.so}; pat=${2:-srbd_mpc_solve}
cuobjdump -res-usage $lib 2>/dev/null | grep -A1 "$pat" | grep -E "Function|REG" | sed 's/^ *//'
for fn in $(cuobjdump -sass $lib | grep "Function :" | grep "$pat" | awk '{print $3}'); do
  echo "== $fn"
  cuobjdump -sass -fun $fn $lib | grep -E "^\s+/\*[0-9a-f]+\*/" | sed -E 's/^\s+\/\*[0-9a-f]+\*\/\s+//' | sed -E 's/^@!?U?P[0-9T]+ //' | awk '{print $1}' | sed 's/;//' > /tmp/sass_ops.txt
  echo "instructions: $(wc -l < /tmp/sass_ops.txt)"
  sort /tmp/sass_ops.txt | uniq -c | sort -rn | head -40
  echo "tensor / TMA mnemonics (UTC*MMA, UTMALDG, HMMA, DMMA, LDGSTS): $(grep -cE 'UTC|UTMA|HMMA|DMMA|LDGSTS' /tmp/sass_ops.txt)"
  echo "local-memory instructions (LDL/STL = spills): $(grep -cE '^LDL|^STL' /tmp/sass_ops.txt)"
done
