# One-line summaries of the kernel-only legs of bench.py for the variants given in $VARIANTS (and $NOCRC_VARIANTS).
export APS_BENCH_QUICK=1 APS_BENCH_PAGES=${APS_BENCH_PAGES:-16384}
show='
import sys,json
for l in sys.stdin:
    if l.startswith("{"):
        d=json.loads(l); print(d["copy_variant"], "value", d["value"], "ms/step", d["ms_per_step"], "frac_measured", round(2*d["value"]/6444.1,4), "put", d["aux"]["put_fused_crc32c_GBps"], "slow_sms", d.get("slow_sms"))
    else: print(l.rstrip()[-200:])
'
for v in ${VARIANTS:-512x4}; do
  APS_COPY_VARIANT=$v timeout 90 python bench.py --steps 300 --warmup 10 --no-cpu-baseline 2>&1 | python -c "$show"
done
for v in ${NOCRC_VARIANTS:-}; do
  APS_BENCH_NOCRC=1 APS_COPY_VARIANT=$v timeout 90 python bench.py --steps 300 --warmup 10 --no-cpu-baseline 2>&1 | python -c "$show"
done
