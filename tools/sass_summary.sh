#!/bin/bash
# Resource usage (registers, spills, shared memory) and SASS opcode histogram of the shipped hot
# kernels, from the built objects (no GPU needed): tools/sass_summary.sh > profiles/<round>_sass_summary.txt
cd "$(dirname "$0")/.."
for obj in mevp transport thermo noalloc; do
  o=nextsimdg_b200/csrc/build/$obj.o
  echo "==== $o: cuobjdump -res-usage"
  cuobjdump -res-usage $o 2>/dev/null | grep -A1 -E "Function .*(mevp_march_kernelILb[01]ELb[01]ELi0|mevp_march_pf_kernel|mevp_fused_kernelILi32ELi8|mevp_pair_kernelILi16ELi16ELi2|transport_march_kernel|transport_pair_march_kernel|transport_stage_kernelILi6ELb1ELi4|transport_stage_kernelILi3ELb1|thermo_kernelILb0ELb0|mass_rows_kernel|mass_final_kernel)" | sed 's/^ *//'
done
cuobjdump -sass nextsimdg_b200/csrc/build/mevp.o > /tmp/_mevp.sass 2>/dev/null
cuobjdump -sass nextsimdg_b200/csrc/build/transport.o > /tmp/_transport.sass 2>/dev/null
hist() { # file, pattern
  awk -v pat="$2" '/Function : /{f=index($0,pat)>0} f' "$1" | grep -oE "^\s+/\*[0-9a-f]+\*/\s+(@!?U?P[0-9] +)?[A-Z0-9_.]+" | awk '{print $NF}' | sed 's/\..*//' | sort | uniq -c | sort -rn | head -24 | awk '{printf "%s %s  ", $2, $1} END {print ""}'
}
echo "==== SASS opcode histogram (whole kernel body: prologue + unrolled copies of the march step)"
echo "mevp_march_kernel<PEER=false, RING=0> (default, single rank):"; hist /tmp/_mevp.sass "mevp_march_kernelILb0ELb0ELi0"
echo "mevp_march_kernel<PEER=true, ld.global.cg, RING=0> (strips, peer stores):"; hist /tmp/_mevp.sass "mevp_march_kernelILb1ELb1ELi0"
echo "transport_pair_march_kernel<6> (dG2, stages 1 + 2 in one launch):"; hist /tmp/_transport.sass "transport_pair_march_kernelILi6"
echo "transport_pair_march_kernel<3> (dG1, the whole Heun step):"; hist /tmp/_transport.sass "transport_pair_march_kernelILi3"
echo "transport_march_kernel<6, HAS_BASE=true>:"; hist /tmp/_transport.sass "transport_march_kernelILi6ELb1"
echo "transport_march_kernel<6, HAS_BASE=false>:"; hist /tmp/_transport.sass "transport_march_kernelILi6ELb0"
echo "transport_stage_kernel<6, base, 4>:"; hist /tmp/_transport.sass "transport_stage_kernelILi6ELb1ELi4"
echo "==== tensor-core mnemonics in the library (expected: none, the path is FP64 streaming):"
cat /tmp/_mevp.sass /tmp/_transport.sass | grep -cE "UTCMMA|HMMA|TCGEN|UTCHMMA" || true
echo "==== bulk-copy engine mnemonics (only in the opt-in experiment arms march_ring = 1 / 3 of mevp.cu):"
grep -oE "UBLKCP[A-Z0-9_.]*|UTMAPF[A-Z0-9_.]*|UBLKPF[A-Z0-9_.]*|UTMALDG[A-Z0-9_.]*" /tmp/_mevp.sass /tmp/_transport.sass | sort | uniq -c
