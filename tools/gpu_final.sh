#!/bin/bash
# One GPU session that produces everything profiles/ and DESIGN.md quote for a revision (round 2).
#   usage (under gpurun): bash tools/gpu_final.sh <tag>      then, in the build container: bash tools/save_final.sh <tag>
TAG=${1:-r2g}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $O/smi.txt 2>&1
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log; tail -2 $O/pytest_gpu.log
python bench.py --steps 10 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; cut -c1-300 $O/bench.json
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; cut -c1-300 $O/bench_ref.json
python tools/quick_bench.py --dtype bf16 --iters 10 > $O/quick_bf16.txt 2>&1; cat $O/quick_bf16.txt
python tools/quick_bench.py --dtype f32 --iters 10 > $O/quick_f32.txt 2>&1; cat $O/quick_f32.txt
python tools/quick_bench.py --dtype bf16 --regime uniform --iters 10 C2_enc C2_dec > $O/quick_bf16_uniform.txt 2>&1; cat $O/quick_bf16_uniform.txt
python bench.py --steps 2 --warmup 3 --e2e-steps 1 --cpu-steps 1 --no-callers > $O/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:msda_ -c 600 --csv --log-file $O/launches.csv \
    python bench.py --steps 2 --warmup 3 --e2e-steps 1 --cpu-steps 1 --no-callers > $O/ncu_launches.log 2>&1
python tools/one_bwd.py C2_enc bf16 structured 3 > $O/plain_one_bwd.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'msda_' -s 4 -c 4 \
    -o $O/prof_enc_bwd python tools/one_bwd.py C2_enc bf16 structured 3 > $O/ncu_full.log 2>&1
python tools/frames_bench.py > $O/frames_1gpu.json 2>/dev/null; cut -c1-200 $O/frames_1gpu.json
python tools/train_bench.py > $O/train_1gpu.json 2>/dev/null; cut -c1-200 $O/train_1gpu.json
python tools/kine_bench.py > $O/kine_bench.json 2>/dev/null; python tools/kine_bench.py --fp32 >> $O/kine_bench.json 2>/dev/null; cut -c1-200 $O/kine_bench.json
python tools/track_phases.py > $O/track_phases.txt 2>&1; cat $O/track_phases.txt
python tools/frame_profile.py --top 20 > $O/frame_breakdown.txt 2>&1; python tools/train_profile.py > $O/train_breakdown.txt 2>&1
ls -la $O
