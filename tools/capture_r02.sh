#!/bin/bash
# Evidence of round 2 (run on the GPU box through gpurun; outputs under gpurun_out/r02/, summarised into profiles/ here).
# Every ncu pass follows a plain run of the same command that exited 0.
set -u
O=gpurun_out/r02; mkdir -p $O
NCU="ncu --set full --clock-control none --import-source on"
run() { echo "== $*" >> $O/log.txt; "$@" >> $O/log.txt 2>&1 || echo "FAILED: $*" >> $O/log.txt; }
# bench lines of the final build
python bench.py --steps 5 --warmup 3 > $O/bench_c4.json 2> $O/bench_c4.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref_c4.json 2> $O/bench_ref_c4.err
python bench.py --workload C5 --steps 3 --warmup 3 --extra "" > $O/bench_c5.json 2> $O/bench_c5.err
# launch list of the headline bench command
run python bench.py --steps 2 --warmup 1 --extra "" --no-cpu-baseline
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches_c4.csv python bench.py --steps 2 --warmup 1 --extra "" --no-cpu-baseline > $O/ncu_launches.log 2>&1
# --set full of the hot kernels (second launch of each: the first is the warm-up)
run python tests/gpu_profile_stage.py C4 eval 1
$NCU -k regex:eval_kernel -s 1 -c 1 -o $O/c4_eval python tests/gpu_profile_stage.py C4 eval 1 >> $O/log.txt 2>&1
run python tests/gpu_profile_stage.py C5 eval 1
$NCU -k regex:eval_kernel -s 1 -c 1 -o $O/c5_eval python tests/gpu_profile_stage.py C5 eval 1 >> $O/log.txt 2>&1
run python tests/gpu_profile_stage.py C4 schur_blocks 1
$NCU -k regex:schur_ -s 4 -c 4 -o $O/c4_schur python tests/gpu_profile_stage.py C4 schur_blocks 1 >> $O/log.txt 2>&1
run python tests/gpu_profile_stage.py C3 cholesky 1
$NCU -k regex:chol_ -s 8 -c 8 -o $O/c3_chol python tests/gpu_profile_stage.py C3 cholesky 1 >> $O/log.txt 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:chol_ -c 400 --csv --log-file $O/launches_c3_chol.csv python tests/gpu_profile_stage.py C3 cholesky 1 >> $O/log.txt 2>&1
run python tests/gpu_profile_stage.py C2 cholesky 1
$NCU -k regex:chol_small -s 1 -c 1 -o $O/c2_chol_small python tests/gpu_profile_stage.py C2 cholesky 1 >> $O/log.txt 2>&1
run python tests/gpu_trace.py C4 3
$NCU -k regex:pcg_blocks -s 1 -c 1 -o $O/c4_pcg python tests/gpu_trace.py C4 3 >> $O/log.txt 2>&1
tail -5 $O/log.txt
