#!/bin/bash
# Runs the text-encoder forward check (tools/te_diag2.py: C forward vs the fp64 torch restatement after 2 and 6 layers) with the shipped
# library (trigger-first PDL, activations read through L2) and with a variant built with plain loads (-DTE_LDCG=0), the combination that
# returned wrong results on some boxes (profiles/r2_te_pdl_race.log).  A box where the variant fails and the shipped library does not
# confirms the fix.  Build the variant first:  see the nvcc line in profiles/r2b_te_race_probe.log
V=$PWD/matcha-tts-etu-upmc-ensam_b200/lib_var/libmb_te_plain.so
nvidia-smi --query-gpu=driver_version,vbios_version --format=csv,noheader
for prec in 0 2; do for s in "3 50" "8 128" "32 128"; do
  echo "shipped  prec=$prec $s: $(timeout 60 python tools/te_diag2.py $s $prec 2>&1 | grep -E 'layers=(2|6)' | sed 's/.*layers=/L/' | tr '\n' ' ')"
  [ -f $V ] && echo "plain-ld prec=$prec $s: $(MB200_LIB=$V timeout 60 python tools/te_diag2.py $s $prec 2>&1 | grep -E 'layers=(2|6)' | sed 's/.*layers=/L/' | tr '\n' ' ')"
done; done
