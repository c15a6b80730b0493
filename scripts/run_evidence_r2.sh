#!/bin/bash
# Evidence pass on one B200: GEMM check + timing, BERT / ResNet kernel breakdowns, memory-bound micro benchmarks, then ONE ncu
# capture of the step-tail kernels (only after the same program exited 0 without ncu). Summaries go to gpurun_out/ as CSV; the
# .ncu-rep stays on the box unless it is small.
mkdir -p gpurun_out
python scripts/gpu_check_gemm.py > gpurun_out/gemm_check_r2.log 2>&1; cp gpurun_out/gemm_timing.json gpurun_out/gemm_timing_r2.json
python scripts/gpu_profile_bert.py > gpurun_out/bert_kernel_breakdown_r2.txt 2>&1
python scripts/gpu_profile_resnet.py > gpurun_out/resnet_kernel_breakdown_r2.txt 2>&1; cp gpurun_out/resnet_timing.json gpurun_out/resnet_timing_r2.json
python scripts/gpu_micro_membound.py > gpurun_out/micro_membound_r2.log 2>&1
python scripts/gpu_ncu_step_kernels.py > gpurun_out/ncu_step_plain.log 2>&1 && \
  ncu --set full --clock-control none --import-source on --profile-from-start off -f -o /tmp/step_kernels_r2 \
      python scripts/gpu_ncu_step_kernels.py > gpurun_out/ncu_step.log 2>&1
if [ -f /tmp/step_kernels_r2.ncu-rep ]; then
  ncu -i /tmp/step_kernels_r2.ncu-rep --page raw --csv > gpurun_out/ncu_step_kernels_r2_raw.csv 2>/dev/null
  ncu -i /tmp/step_kernels_r2.ncu-rep --page details --csv > gpurun_out/ncu_step_kernels_r2_details.csv 2>/dev/null
  sz=$(stat -c %s /tmp/step_kernels_r2.ncu-rep); echo "ncu-rep bytes: $sz" >> gpurun_out/ncu_step.log
  if [ "$sz" -lt 30000000 ]; then cp /tmp/step_kernels_r2.ncu-rep gpurun_out/; fi
fi
du -sh gpurun_out; for f in gpurun_out/*.log; do echo "== $f"; tail -n 3 "$f"; done
