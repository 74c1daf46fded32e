# Sweeps the launch shapes of the copy kernel on one GPU (kernel-only legs of bench.py).
export APS_BENCH_QUICK=1 APS_BENCH_PAGES=${APS_BENCH_PAGES:-16384}
show='
import sys,json
for l in sys.stdin:
    if l.startswith("{"):
        d=json.loads(l); print(d["copy_variant"], "value", d["value"], "ms/step", d["ms_per_step"], "copy_ms", d["roofline"]["avg_launch_ms"], "frac", d["roofline"]["frac"], "put", d["aux"]["put_fused_crc32c_GBps"], "launches", d["gpu_launches"], "slow_sms", d.get("slow_sms"))
    else: print(l.rstrip())
'
for v in ${VARIANTS:-512x4 1024x4 512x8}; do
  APS_COPY_VARIANT=$v python bench.py --steps 300 --warmup 10 --no-cpu-baseline 2>&1 | python -c "$show"
done
for v in ${NOCRC_VARIANTS:-512x4}; do
  APS_BENCH_NOCRC=1 APS_COPY_VARIANT=$v python bench.py --steps 300 --warmup 10 --no-cpu-baseline 2>&1 | python -c "$show"
done
