#!/bin/bash
# Round routine on the GPU box: tests, smoke, bench, ncu launch list + full capture of the top kernel.
set -u
TAG=${1:-r01}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" 
timeout 300 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
SHORT="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-path-kernels"
timeout 300 $SHORT > gpurun_out/${TAG}_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $SHORT > gpurun_out/${TAG}_ncu1.log 2>&1; echo "ncu list rc=$?"
timeout 300 $SHORT > gpurun_out/${TAG}_plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:corr_ -s 3 -c 2 \
    -f -o gpurun_out/${TAG}_corr $SHORT > gpurun_out/${TAG}_ncu2.log 2>&1; echo "ncu full rc=$?"
# duration + DRAM bytes of every other path kernel (CorrAcc, CorrSubsel, reorder, Beamform, BeamformSum)
timeout 300 python scripts/prof_path.py > gpurun_out/${TAG}_path_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    -k regex:"acc_kernel|subsel_kernel|reorder_tile|beamform_kernel|beam_integrate" -c 40 --csv \
    --log-file gpurun_out/${TAG}_path_kernels.csv python scripts/prof_path.py > gpurun_out/${TAG}_ncu3.log 2>&1; echo "ncu path rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:beamform_kernel -s 2 -c 1 \
    -f -o gpurun_out/${TAG}_beam python scripts/prof_beam.py > gpurun_out/${TAG}_ncu4.log 2>&1; echo "ncu beam rc=$?"
# offline upchannelise kernels: duration + DRAM bytes per launch
timeout 300 python scripts/prof_upchan.py > gpurun_out/${TAG}_upchan_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none \
    -k regex:"upchan_" -c 12 --csv \
    --log-file gpurun_out/${TAG}_upchan_kernels.csv python scripts/prof_upchan.py > gpurun_out/${TAG}_ncu5.log 2>&1; echo "ncu upchan rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:upchan_beamform -s 1 -c 1 \
    -f -o gpurun_out/${TAG}_upchan_bf python scripts/prof_upchan.py > gpurun_out/${TAG}_ncu6.log 2>&1; echo "ncu upchan full rc=$?"
# the reference arm (CPU formulation on the host cores), as the driver runs it
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; echo "reference arm rc=$?"
