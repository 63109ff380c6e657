# Round-1 evidence run (one B200): bench lines, ncu launch list of the bench command, one
# `ncu --set full` capture per hot kernel at the c2 launch size.  Outputs under gpurun_out/.
set -x
T=${1:-r1c}
python bench.py --steps 10 --warmup 3 > gpurun_out/${T}_bench_c2_n1.json 2> gpurun_out/${T}_bench_c2_n1.err
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/${T}_plain_c2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_ncu_launches_c2.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/${T}_ncu_launch.log 2>&1
python tools/time_fused.py 2048 512 30 > gpurun_out/${T}_plain_full.log 2>&1 && for k in stft_fwd istft_inv ola_gather istft_ola; do ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 1 -o gpurun_out/${T}_full_$k python tools/time_fused.py 2048 512 30 > gpurun_out/${T}_ncu_$k.log 2>&1; done
cat gpurun_out/${T}_plain_full.log
python bench.py --steps 5 --warmup 3 --workload c3 --no-cpu > gpurun_out/${T}_bench_c3_n1.json 2> gpurun_out/${T}_bench_c3_n1.err
python bench.py --steps 5 --warmup 3 --workload c4 --no-cpu > gpurun_out/${T}_bench_c4_n1.json 2> gpurun_out/${T}_bench_c4_n1.err
python bench.py --steps 5 --warmup 3 --workload c5 --no-cpu > gpurun_out/${T}_bench_c5_n1.json 2> gpurun_out/${T}_bench_c5_n1.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err
tail -c 1200 gpurun_out/${T}_bench_c2_n1.json
