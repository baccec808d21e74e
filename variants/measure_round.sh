set -x
python -m pytest tests -m gpu -q 2>&1 | tail -3
python bench.py > gpurun_out/r1_f_bench.json 2> gpurun_out/r1_f_bench.err
python bench.py --impl reference > gpurun_out/r1_f_bench_reference.json 2> gpurun_out/r1_f_bench_reference.err
python bench.py --steps 50 --warmup 5 > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_f_launches.csv python bench.py --steps 50 --warmup 5 > gpurun_out/ncu_launches.log 2>&1
python variants/few_frames.py && ncu --set full --import-source on --clock-control none -k regex:"ocean_(rows|cols_tma)_kernel" -s 4 -c 4 -o gpurun_out/r1_f_full -f python variants/few_frames.py > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/r1_f_full.ncu-rep --page raw --csv > gpurun_out/r1_f_full_raw.csv 2>/dev/null
python variants/ncu_summary.py gpurun_out/r1_f_full_raw.csv > gpurun_out/r1_f_ncu_full_summary.txt
(python variants/config_rates.py; python variants/ensemble_rate.py | cut -c1-110; python variants/time_mips.py) > gpurun_out/r1_f_configs.txt 2>&1
tail -3 gpurun_out/r1_f_bench.json | cut -c1-600
