#!/bin/bash
# round-2 ncu evidence: MAS kernel --set full on the shipped kernel (BASELINE configs[1]); decoder DRAM bytes per solve; top decoder kernel page
cd $GRAFT_REPO_ROOT
python tools/mas_one.py 1024 400 2000 3 > gpurun_out/r2_ncu2_mas_plain.log 2>&1 || { echo "mas plain failed"; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mas_kernel -s 2 -c 1 -f -o gpurun_out/r2_mas_full \
    python tools/mas_one.py 1024 400 2000 3 > gpurun_out/r2_ncu2_mas.log 2>&1
echo "mas ncu rc=$?"
ncu -i gpurun_out/r2_mas_full.ncu-rep --page raw --csv > gpurun_out/r2_mas_ncu_raw.csv 2>/dev/null
python tools/decoder_one.py 32 800 10 fp16 1 > gpurun_out/r2_ncu2_dec_plain.log 2>&1 || { echo "dec plain failed"; exit 1; }
for cc in none all; do
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --cache-control $cc --csv \
    --log-file gpurun_out/r2_decoder_fp16_dram_cache_$cc.csv python tools/decoder_one.py 32 800 10 fp16 1 > gpurun_out/r2_ncu2_dec_$cc.log 2>&1
echo "dec dram ($cc) rc=$?"; wc -l gpurun_out/r2_decoder_fp16_dram_cache_$cc.csv
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ffn_fused -s 12 -c 1 -f -o gpurun_out/r2_ffn_outproj_full \
    python tools/decoder_one.py 32 800 1 fp16 1 > gpurun_out/r2_ncu2_ffn.log 2>&1
echo "ffn ncu rc=$?"
ncu -i gpurun_out/r2_ffn_outproj_full.ncu-rep --page raw --csv > gpurun_out/r2_ffn_outproj_ncu_raw.csv 2>/dev/null
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attention_tc -s 12 -c 1 -f -o gpurun_out/r2_attn_full \
    python tools/decoder_one.py 32 800 1 fp16 1 > gpurun_out/r2_ncu2_attn.log 2>&1
echo "attn ncu rc=$?"
ncu -i gpurun_out/r2_attn_full.ncu-rep --page raw --csv > gpurun_out/r2_attn_ncu_raw.csv 2>/dev/null
ls -la gpurun_out/ | tail -20
