# Round profile on one B200: plain bench, ncu launch list of the same command, one full capture of the
# copy kernel.  Outputs land in gpurun_out/ (copy the summaries into profiles/ afterwards).
tag=${1:-r1}
mkdir -p gpurun_out
python bench.py --steps 500 --warmup 10 > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
tail -c 2600 gpurun_out/bench_$tag.json
export APS_BENCH_QUICK=1 APS_BENCH_PAGES=8192
cmd="python bench.py --steps 6 --warmup 3 --no-cpu-baseline"
$cmd > gpurun_out/plain_$tag.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$tag.csv $cmd > gpurun_out/ncu_list_$tag.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:copy_kernel -s 41 -c 1 -o gpurun_out/prof_$tag $cmd > gpurun_out/ncu_full_$tag.log 2>&1
tail -n 2 gpurun_out/ncu_full_$tag.log
