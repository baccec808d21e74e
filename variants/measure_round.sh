# Round measurements on one B200 box (run under gpurun from the repo root):  bash variants/measure_round.sh r2
T=${1:-r2}
set -x
python -m pytest tests -m gpu -q 2>&1 | tail -3 > gpurun_out/${T}_gpu_tests.txt
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err
python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench_k20.json 2>> gpurun_out/${T}_bench.err
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err
B="python bench.py --steps 50 --warmup 5 --no-ensemble --no-cpu-baseline"
$B > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv $B > gpurun_out/ncu_launches.log 2>&1
$B > /dev/null 2>&1 && ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --cache-control none --clock-control none -k regex:"ocean_(rows|cols_tma)_kernel" -s 60 -c 120 --csv --log-file gpurun_out/${T}_warm_traffic.csv $B > gpurun_out/ncu_warm.log 2>&1
python variants/few_frames.py && ncu --set full --import-source on --clock-control none -k regex:"ocean_(rows|cols_tma)_kernel" -s 4 -c 4 -o gpurun_out/${T}_full -f python variants/few_frames.py > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/${T}_full.ncu-rep --page raw --csv > gpurun_out/${T}_full_raw.csv 2>/dev/null
python variants/ncu_summary.py gpurun_out/${T}_full_raw.csv > gpurun_out/${T}_ncu_full_summary.txt
(python variants/config_rates.py; python variants/ensemble_rate.py | cut -c1-110; python variants/time_mips.py) > gpurun_out/${T}_configs.txt 2>&1
python variants/trace_frame.py > gpurun_out/${T}_phase_trace_two_kernel.txt 2>&1
cut -c1-400 gpurun_out/${T}_bench.json
