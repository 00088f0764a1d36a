# third ncu call of round 2: the JSON encode kernels and the AB-test kernel (reduced on the box like the others)
set -x
python tools/prof_workload.py codec > gpurun_out/prof_plain_codec2.log 2>&1 && \
ncu --set full --clock-control none -k regex:"json_fmt|abtest" --launch-skip 0 --launch-count 10 -f -o gpurun_out/prof_r02_enc python tools/prof_workload.py codec > gpurun_out/ncu_enc.log 2>&1
python tools/ncu_summary.py gpurun_out/prof_r02_enc.ncu-rep > gpurun_out/prof_r02_enc_summary.txt 2>&1
rm -f gpurun_out/prof_r02_enc.ncu-rep
