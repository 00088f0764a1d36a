# ncu captures of round 2 (one gpurun call; every profiled command first runs plain and must exit 0).
# The .ncu-rep files are too large to bring back: they are reduced on the box to the raw metric table (CSV), the
# metric summary of tools/ncu_summary.py and the hottest instructions of tools/ncu_hot.py, then removed.
set -x
reduce() {  # $1 = report stem
  ncu -i gpurun_out/$1.ncu-rep --page raw --csv > gpurun_out/$1_raw.csv 2>/dev/null
  python tools/ncu_summary.py gpurun_out/$1.ncu-rep > gpurun_out/$1_summary.txt 2>&1
  python tools/ncu_hot.py gpurun_out/$1.ncu-rep > gpurun_out/$1_hot.txt 2>&1
  rm -f gpurun_out/$1.ncu-rep
}
python tools/prof_workload.py c2 > gpurun_out/prof_plain_c2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"pass1_gram_umma|project_rows|chain4|combiner_mean" --launch-skip 16 --launch-count 4 -f -o gpurun_out/prof_r02_c2 python tools/prof_workload.py c2 > gpurun_out/ncu_c2.log 2>&1
reduce prof_r02_c2
python tools/prof_workload.py c5 > gpurun_out/prof_plain_c5.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"pass1_gram_umma|project_rows|chain4|subspace_topk" --launch-skip 11 --launch-count 4 -f -o gpurun_out/prof_r02_c5 python tools/prof_workload.py c5 > gpurun_out/ncu_c5.log 2>&1
reduce prof_r02_c5
python tools/prof_workload.py c5 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"jacobi_cluster" --launch-skip 0 --launch-count 1 -f -o gpurun_out/prof_r02_jacobi python tools/prof_workload.py c5 > gpurun_out/ncu_jac.log 2>&1
reduce prof_r02_jacobi
python tools/prof_workload.py codec > gpurun_out/prof_plain_codec.log 2>&1 && \
ncu --set full --clock-control none -k regex:"pack_|unpack_|json_|combiner_mean" --launch-skip 0 --launch-count 30 -f -o gpurun_out/prof_r02_codec python tools/prof_workload.py codec > gpurun_out/ncu_codec.log 2>&1
reduce prof_r02_codec
python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-blocks > gpurun_out/r02_launch_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-blocks > gpurun_out/ncu_launches.log 2>&1
du -sh gpurun_out
