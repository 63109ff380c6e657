# Round-2 evidence run (one B200): bench lines, the ncu launch list of the bench command, one `ncu --set full`
# capture per hot kernel at the c2 / c4 launch sizes (exported to raw CSV on the box; the .ncu-rep files stay there).
# Outputs under gpurun_out/.   usage: bash tools/run_prof.sh [tag]
set -x
T=${1:-r2}
python bench.py --steps 10 --warmup 3 > gpurun_out/${T}_bench_c2_n1.json 2> gpurun_out/${T}_bench_c2_n1.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_reference_n1.json 2> gpurun_out/${T}_bench_reference_n1.err
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-legs > gpurun_out/${T}_plain_c2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_ncu_launches_c2.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-legs > gpurun_out/${T}_ncu_launch.log 2>&1
for cfg in "2048 512" "4096 1024"; do
  set -- $cfg
  python tools/time_fused.py $1 $2 30 > gpurun_out/${T}_plain_full_$1.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'stft_fwd|istft_inv|ola_gather|istft_ola' -s 8 -c 26 -f -o /tmp/${T}_full_$1 python tools/time_fused.py $1 $2 30 > gpurun_out/${T}_ncu_$1.log 2>&1
  ncu -i /tmp/${T}_full_$1.ncu-rep --page raw --csv > gpurun_out/${T}_ncu_full_$1.raw.csv
  cat gpurun_out/${T}_plain_full_$1.log
done
tail -c 600 gpurun_out/${T}_bench_c2_n1.json
