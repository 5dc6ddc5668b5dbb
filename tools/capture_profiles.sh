#!/usr/bin/env bash
# Run under gpurun (one GPU): launch list of a 2^28 x 3 build and one `ncu --set full` capture per hot kernel.
set -u
python tools/profile_build.py --log2n 28 > gpurun_out/plain28.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches28.csv python tools/profile_build.py --log2n 28 > gpurun_out/ncu28.log 2>&1
cat gpurun_out/plain28.log
python tools/profile_build.py --log2n 26 > gpurun_out/plain26.log 2>&1 || exit 1
for k in k_partition:2 k_subtree:0 k_medium:1 k_bracket:2 k_hist:13 k_init:0; do
    name=${k%%:*}; skip=${k##*:}
    ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:$name -s $skip -c 1 \
        -o gpurun_out/prof_$name -f python tools/profile_build.py --log2n 26 > gpurun_out/ncu_$name.log 2>&1
done
cat gpurun_out/plain26.log
