#!/bin/bash
# Copy the judged outputs of one tools/gpu_round.sh session from gpurun_out/<tag>/ (scratch) into profiles/ (tracked).
TAG=${1:?usage: tools/save_profiles.sh <tag>}
O=gpurun_out/$TAG
cp $O/bench.json profiles/${TAG}_bench.json
cp $O/bench_ref.json profiles/${TAG}_bench_ref.json
cp $O/quick_bf16.txt profiles/${TAG}_quick_bf16.txt
cp $O/quick_f32.txt profiles/${TAG}_quick_f32.txt
cp $O/quick_f32_uniform.txt profiles/${TAG}_quick_f32_uniform.txt
cp $O/ref_kernel_timing.txt profiles/${TAG}_reference_kernel_and_stack_timing.txt
cp $O/launches.csv profiles/${TAG}_launches_bench_steps2_msda_kernels.csv
python tools/ncu_metrics.py $O/prof_enc_bf16.ncu-rep > profiles/${TAG}_ncu_full_C2enc_bf16.txt
[ -f gpurun_out/frame_graph_timing.json ] && cp gpurun_out/frame_graph_timing.json profiles/${TAG}_frame_graph_timing.json
[ -f gpurun_out/transformer_stack_timing.json ] && cp gpurun_out/transformer_stack_timing.json profiles/${TAG}_transformer_stack_timing.json
ls profiles/${TAG}_*
