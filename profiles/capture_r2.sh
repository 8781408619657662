#!/bin/bash
# Round-2 evidence, run on the GPU box (gpurun): per workload an ncu launch list of the bench command
# (gpu__time_duration per launch) and one `--set full` capture of every dominant kernel.  Each bench
# command has already exited 0 without ncu before it is run under it.  Outputs go to gpurun_out/.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
NCU="ncu --clock-control none"
# launch lists: this library's kernels only (the generators of the synthetic inputs are torch kernels
# outside the timed region and would fill the list)
for wl in afsl_c3 tiresias_c1 ffdl_c2 munkres_c4 alloc_place; do
  python bench.py --workload $wl --only-headline --steps 2 --warmup 3 --e2e-steps 1 --no-cpu > gpurun_out/r2_plain_$wl.json 2> gpurun_out/r2_plain_$wl.err || { echo "plain $wl failed"; continue; }
  $NCU --metrics gpu__time_duration.sum -k "regex:^(afsl|tiresias|ffdl|fill|munkres|place|curves|tick|compare)" -c 400 --csv --log-file gpurun_out/r2_launches_$wl.csv \
      python bench.py --workload $wl --only-headline --steps 2 --warmup 3 --e2e-steps 1 --no-cpu > /dev/null 2>&1
done
cap() {  # name, kernel regex, workload, launch-skip
  $NCU --set full --import-source on -k regex:$2 --launch-skip $4 -c 1 -f -o gpurun_out/r2_$1 \
      python bench.py --workload $3 --only-headline --steps 2 --warmup 3 --e2e-steps 1 --no-cpu > gpurun_out/r2_$1.log 2>&1
}
cap afsl afsl_warp_kernel afsl_c3 3
cap tiresias tiresias_kernel tiresias_c1 3
cap ffdl_dp ffdl_dp_kernel ffdl_c2 3
cap ffdl_bt ffdl_backtrack_kernel ffdl_c2 3
cap munkres munkres_kernel munkres_c4 4   # launch 4 = the uint8 pass of the second step's ladder
cap place place_kernel alloc_place 3
cap fill fill_kernel alloc_place 3
ls -la gpurun_out | tail -30
