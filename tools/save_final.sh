#!/bin/bash
# Copy the judged outputs of tools/gpu_final.sh from gpurun_out/<tag>/ (scratch) into profiles/ (tracked) and refresh
# profiles/roofline_inputs.json (keyed by the hash of kinet_b200/csrc).
TAG=${1:?usage: tools/save_final.sh <tag>}
O=gpurun_out/$TAG
for f in bench.json bench_ref.json quick_bf16.txt quick_f32.txt quick_bf16_uniform.txt kine_bench.json track_phases.txt; do cp $O/$f profiles/${TAG}_$f; done
cp $O/launches.csv profiles/${TAG}_launches_bench_steps2_msda_kernels.csv
cut -c1-170 $O/frame_breakdown.txt > profiles/${TAG}_frame_kernel_breakdown.txt
cut -c1-170 $O/train_breakdown.txt > profiles/${TAG}_train_step_kernel_breakdown.txt
cat $O/frames_1gpu.json $O/train_1gpu.json > profiles/${TAG}_callers_tools.json
python tools/ncu_metrics.py $O/prof_enc_bwd.ncu-rep > profiles/${TAG}_ncu_full_C2enc_bf16_all_kernels.txt
python tools/roofline_inputs.py $O/prof_enc_bwd.ncu-rep profiles/${TAG}_ncu_full_C2enc_bf16_backward.txt
python tools/sass_excerpts.py > profiles/${TAG}_sass_excerpts.txt 2>&1
ls profiles/${TAG}_* profiles/roofline_inputs.json
