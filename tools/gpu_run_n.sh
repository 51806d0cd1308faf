#!/bin/bash
# usage: tools/gpu_run_n.sh N [steps]   -- multi-GPU bench through torchrun (+ the 2-rank NCCL test when N == 2)
N=${1:-2}; STEPS=${2:-3}
mkdir -p gpurun_out
if [ "$N" = "2" ]; then
  python -m pytest tests/test_gpu_multi.py -m gpu -q -x > gpurun_out/pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -5 gpurun_out/pytest_multi.log
fi
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps $STEPS --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
tail -5 gpurun_out/bench_n$N.err
python - <<PY
import json
l=json.loads([x for x in open('gpurun_out/bench_n$N.json') if x.startswith('{')][-1])
print({k:l.get(k) for k in ('value','ms_per_step','n_gpus','state_broadcast_s','state_broadcast_bytes','sharded_vs_unsharded_max_rel_dev','comm_nranks_seen')})
print('e2e', l['e2e']); print('clocks', l['clocks_all_ranks'])
print('roof', {k:l['roofline'].get(k) for k in ('frac','frac_of_sustained_peak','step_frac','ms_per_launch')})
ex=l.get('extras',{})
for k in ('config4a','config4b','config5'):
    print(k, json.dumps(ex.get(k))[:1200])
print('mll', json.dumps(l.get('mll'))[:800])
PY
