#!/bin/bash
# One GPU call that produces the round's evidence files under gpurun_out/ (copied into profiles/ afterwards):
#   bench line, ncu launch list of the same command, ncu --set full captures of the dominant kernels.
# Usage (on the GPU box): bash tools/profile_round.sh r02
R=${1:-rXX}
mkdir -p gpurun_out
python bench.py --steps 5 --warmup 3 > gpurun_out/${R}_bench_1gpu.json 2> gpurun_out/${R}_bench_1gpu.err || tail -5 gpurun_out/${R}_bench_1gpu.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${R}_launches_bench.csv \
    python bench.py --steps 2 --warmup 3 > gpurun_out/${R}_launches_bench.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:self_apply_fixed --launch-skip 3 -c 1 -o gpurun_out/${R}_k1 \
    python bench.py --steps 2 --warmup 3 --no-legs > gpurun_out/${R}_k1_ncu.log 2>&1
ncu -i gpurun_out/${R}_k1.ncu-rep --page raw --csv > gpurun_out/${R}_k1_raw.csv 2>/dev/null
ncu --set full --clock-control none --import-source on -k regex:"soup_sync_kernel|soup_deferred|soup_lists" --launch-skip 6 -c 3 -o gpurun_out/${R}_soup10m \
    python tools/prof_soup.py 10000000 4 0 > gpurun_out/${R}_soup10m_ncu.log 2>&1
ncu -i gpurun_out/${R}_soup10m.ncu-rep --page raw --csv > gpurun_out/${R}_soup10m_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/${R}_k1_raw.csv
python tools/ncu_summary.py gpurun_out/${R}_soup10m_raw.csv
rm -f gpurun_out/${R}_k1.ncu-rep gpurun_out/${R}_soup10m.ncu-rep
