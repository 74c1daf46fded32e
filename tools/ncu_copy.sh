# ncu captures of the copy kernel (one GPU).  Usage: bash tools/ncu_copy.sh <tag> ; extra env passes through.
tag=${1:-r1}
export APS_BENCH_QUICK=1 APS_BENCH_PAGES=${APS_BENCH_PAGES:-8192}
cmd="python bench.py --steps 6 --warmup 3 --no-cpu-baseline"
$cmd > gpurun_out/plain_$tag.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:copy_kernel -s 40 -c 1 -o gpurun_out/prof_${tag} $cmd > gpurun_out/ncu_$tag.log 2>&1
tail -n 2 gpurun_out/ncu_$tag.log
