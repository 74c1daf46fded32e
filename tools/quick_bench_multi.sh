# Kernel-only sharded bench on N GPUs for the copy-kernel variants in $VARIANTS.  Usage: N=2 VARIANTS="512x4 512x8" bash tools/quick_bench_multi.sh
export APS_BENCH_QUICK=1
N=${N:-2}
port=29600
for v in ${VARIANTS:-512x4}; do
  port=$((port+1))
  APS_COPY_VARIANT=$v timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N --steps 200 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); r=d['roofline']; print(d['copy_variant'], 'n', d['n_gpus'], 'value', d['value'], 'ms/step', d['ms_per_step'], 'nvlink', r.get('achieved'), 'frac', r.get('frac'), 'remote', r.get('remote_fraction'))
"
done
