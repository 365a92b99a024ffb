#!/bin/bash
# One GPU session: parity tests, contract bench, per-shape timings, ncu launch list + full capture.
# usage (under gpurun): bash tools/gpu_round.sh <tag>
TAG=${1:-r1}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $O/smi.txt 2>&1
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
python bench.py --steps 10 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench exit $?"; cat $O/bench.json
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; cat $O/bench_ref.json
python tools/quick_bench.py --dtype f32 --iters 10 > $O/quick_f32.txt 2>&1; cat $O/quick_f32.txt
python tools/quick_bench.py --dtype bf16 --iters 10 > $O/quick_bf16.txt 2>&1; cat $O/quick_bf16.txt
python tools/quick_bench.py --dtype f32 --regime uniform --iters 10 C2_enc C2_dec > $O/quick_f32_uniform.txt 2>&1; cat $O/quick_f32_uniform.txt
python -m pytest tests/test_full_size_gpu.py tests/test_transformer_gpu.py -m gpu -q -s -k "timing or speed" > $O/ref_kernel_timing.txt 2>&1; tail -30 $O/ref_kernel_timing.txt
python bench.py --steps 2 --warmup 3 --e2e-steps 1 --cpu-steps 1 > $O/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:msda_ -c 600 --csv --log-file $O/launches.csv \
    python bench.py --steps 2 --warmup 3 --e2e-steps 1 --cpu-steps 1 > $O/ncu_launches.log 2>&1
python tools/quick_bench.py C2_enc --dtype bf16 --iters 2 > $O/plain_qb.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'msda_(forward|backward)_kernel' -s 3 -c 4 \
    -o $O/prof_enc_bf16 python tools/quick_bench.py C2_enc --dtype bf16 --iters 2 > $O/ncu_full.log 2>&1
python tools/layer_bench.py > $O/layer_bench.txt 2>&1; cat $O/layer_bench.txt
python tools/frames_bench.py > $O/frames_1gpu.json 2>/dev/null; python tools/frames_bench.py --varying-queries >> $O/frames_1gpu.json 2>/dev/null; cat $O/frames_1gpu.json
python tools/train_bench.py > $O/train_1gpu.json 2>/dev/null; cat $O/train_1gpu.json
python tools/frame_profile.py --top 20 > $O/frame_breakdown.txt 2>&1; python tools/frame_profile.py --train --top 30 > $O/train_breakdown.txt 2>&1
ls -la $O
