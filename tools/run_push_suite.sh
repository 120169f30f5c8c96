#!/bin/bash
# Push exchange (producer kernels store into the peers' mailboxes) vs the put kernel: parity at NG ranks, then the
# bench and cfg4 with GSPLAT_B200_EXCHANGE_PUSH=1/0. Outputs under gpurun_out/r02s_*.
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
NG=$(nvidia-smi -L | wc -l)
echo "GPUs: $NG"
timeout 900 $TR --nproc-per-node $NG --master-port 29511 tools/dist_check.py > gpurun_out/r02s_dist_check$NG.log 2>&1
echo "dist_check rc=$?"; grep "DIST_CHECK\|MISMATCH\|Error\|error" gpurun_out/r02s_dist_check$NG.log | head -5 | cut -c1-400
for rep in a b; do
for push in 1 0; do
  GSPLAT_B200_EXCHANGE_PUSH=$push timeout 300 $TR --nproc-per-node $NG --master-port 2952$push bench.py --gpus $NG --steps 50 --warmup 5 \
      > gpurun_out/r02s_bench${NG}_push${push}_$rep.json 2> gpurun_out/r02s_bench${NG}_push${push}_$rep.err
  echo "bench N=$NG push=$push rc=$?"
done
done
for push in 1 0; do
  GSPLAT_B200_EXCHANGE_PUSH=$push CFG4_PACKED=0 CFG4_STEPS=10 timeout 300 $TR --nproc-per-node $NG --master-port 2954$push tools/bench_cfg4.py \
      > gpurun_out/r02s_cfg4_w${NG}_push$push.log 2>&1
  echo "cfg4 world=$NG push=$push rc=$?"
done
grep -h CFG4 gpurun_out/r02s_cfg4_w*.log | cut -c1-400
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02s_bench*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d.get('parity',{}).get('img_max_err'), d.get('parity',{}).get('grad_max_rel'))
    except Exception as e: print(f, 'ERR', e)
PY
