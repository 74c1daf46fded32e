# ncu captures of the kernels other than the headline get: the fused put and the K1 table kernels (one GPU).
# Usage: bash tools/ncu_extra.sh put | table     (each: the plain command first, then ONE ncu run of the same command)
what=${1:-put}
mkdir -p gpurun_out
if [ "$what" = put ]; then
  export APS_BENCH_QUICK=1 APS_BENCH_PAGES=8192
  cmd="python bench.py --steps 6 --warmup 3 --no-cpu-baseline"
  $cmd > gpurun_out/plain_put.log 2>&1 && \
  ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:copy_kernelILb1ELi512ELi4ELb0ELb1 -s 5 -c 1 -o gpurun_out/prof_put $cmd > gpurun_out/ncu_put.log 2>&1
  tail -n 2 gpurun_out/ncu_put.log
else
  cmd="python tools/bench_table.py --ids 2097152 --batch 1048576 --reps 2"
  $cmd > gpurun_out/plain_table.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:"lookup_kernel|delete_kernel" -s 2 -c 2 -o gpurun_out/prof_table $cmd > gpurun_out/ncu_table.log 2>&1
  tail -n 2 gpurun_out/ncu_table.log
fi
