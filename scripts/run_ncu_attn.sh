#!/bin/bash
# ncu capture of the single-tile attention kernels (forward dense form + backward), summaries as CSV
mkdir -p gpurun_out
python scripts/gpu_ncu_attn128.py > gpurun_out/ncu_attn_plain.log 2>&1 && \
  ncu --set full --clock-control none --import-source on --kernel-name regex:attention --launch-skip 4 --launch-count 2 -f -o /tmp/attn_r2 \
      python scripts/gpu_ncu_attn128.py > gpurun_out/ncu_attn.log 2>&1
if [ -f /tmp/attn_r2.ncu-rep ]; then
  ncu -i /tmp/attn_r2.ncu-rep --page raw --csv > gpurun_out/ncu_attn_r2_raw.csv 2>/dev/null
  ncu -i /tmp/attn_r2.ncu-rep --page details --csv > gpurun_out/ncu_attn_r2_details.csv 2>/dev/null
  ncu -i /tmp/attn_r2.ncu-rep --page source --csv --kernel-name regex:attention_fwd > gpurun_out/ncu_attn_fwd_r2_source.csv 2>/dev/null
  ncu -i /tmp/attn_r2.ncu-rep --page source --csv --kernel-name regex:attention_bwd > gpurun_out/ncu_attn_bwd_r2_source.csv 2>/dev/null
fi
tail -3 gpurun_out/ncu_attn.log; ls -la gpurun_out | head
