# second ncu call of round 2: the d = 64 tcgen05 Gram kernel after the truncation split (tensor-pipe counters, DRAM
# bytes) and the launch list of the final bench command.  Reports are reduced on the box (see ncu_call_r02.sh).
set -x
reduce() {
  ncu -i gpurun_out/$1.ncu-rep --page raw --csv > gpurun_out/$1_raw.csv 2>/dev/null
  python tools/ncu_summary.py gpurun_out/$1.ncu-rep > gpurun_out/$1_summary.txt 2>&1
  python tools/ncu_hot.py gpurun_out/$1.ncu-rep > gpurun_out/$1_hot.txt 2>&1
  rm -f gpurun_out/$1.ncu-rep
}
python tools/prof_workload.py c2 > gpurun_out/prof_plain_c2b.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"pass1_gram_umma" --launch-skip 4 --launch-count 1 -f -o gpurun_out/prof_r02_pass1 python tools/prof_workload.py c2 > gpurun_out/ncu_pass1.log 2>&1
reduce prof_r02_pass1
python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-blocks > gpurun_out/r02_launch_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-blocks > gpurun_out/ncu_launches.log 2>&1
du -sh gpurun_out
