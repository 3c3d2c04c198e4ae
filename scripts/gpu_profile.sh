#!/bin/bash
# ncu evidence on the GPU box: launch list of the bench command, full captures of the correlator and the beamformer.
set -u
TAG=${1:-r02}
mkdir -p gpurun_out
SHORT="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-path-kernels --no-xengine-full"
timeout 300 $SHORT > gpurun_out/${TAG}_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $SHORT > gpurun_out/${TAG}_ncu1.log 2>&1; echo "ncu list rc=$?"
timeout 300 $SHORT > gpurun_out/${TAG}_plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:corr_swz -s 4 -c 2 \
    -f -o gpurun_out/${TAG}_corr $SHORT > gpurun_out/${TAG}_ncu2.log 2>&1; echo "ncu corr rc=$?"
timeout 300 python scripts/prof_beam.py > gpurun_out/${TAG}_beam_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:beamform_tc -s 2 -c 2 \
    -f -o gpurun_out/${TAG}_beam python scripts/prof_beam.py > gpurun_out/${TAG}_ncu4.log 2>&1; echo "ncu beam rc=$?"
timeout 300 python scripts/prof_path.py > gpurun_out/${TAG}_path_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    -k regex:"acc_kernel|subsel_kernel|reorder_tile|beamform|beam_integrate|beam_weight" -c 60 --csv \
    --log-file gpurun_out/${TAG}_path_kernels.csv python scripts/prof_path.py > gpurun_out/${TAG}_ncu3.log 2>&1; echo "ncu path rc=$?"
