#!/bin/bash
# round-2 profiles: launch list of the bench command, then --set full on the row kernels (C2) and on the C3 update
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD="python bench.py --steps 24 --warmup 8 --quick"
$CMD > gpurun_out/r2_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 400 --csv --log-file gpurun_out/r2_launches_cold.csv $CMD > gpurun_out/r2_ncu1.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/r2_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_chunk_reduce|k_gather|k_radix_pass|k_prepare" -s 60 -c 12 -o gpurun_out/r2_prof_c2 -f $CMD > gpurun_out/r2_ncu2.log 2>&1
echo "full c2 rc=$?"
CMD3="python tools/phase_bench.py C3_adam --reps 6"
$CMD3 > gpurun_out/r2_plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_chunk_reduce" -s 4 -c 3 -o gpurun_out/r2_prof_c3 -f $CMD3 > gpurun_out/r2_ncu3.log 2>&1
echo "full c3 rc=$?"
ls -la gpurun_out/*.ncu-rep gpurun_out/r2_launches_cold.csv
# bit-exact order mode kernels (C2)
CMD4="python tools/phase_bench.py C2_ctr_sgd --det-only"
$CMD4 > gpurun_out/r2_plain4.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_long_segment_det|k_segment_det" -c 4 -o gpurun_out/r2_prof_det -f $CMD4 > gpurun_out/r2_ncu4.log 2>&1
echo "full det rc=$?"
