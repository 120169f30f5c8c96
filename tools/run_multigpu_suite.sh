#!/bin/bash
# One 8-GPU box, everything the multi-GPU rows need (VERDICT r1 "next" 1d, 2, 7): parity at 8 ranks, the bench at
# 1/2/4/8 with the parity field, split vs unsplit exchange, cfg4 (6 M Gaussians) unpacked and packed at 1/2/4/8 ranks,
# host<->device bandwidth per rank. Outputs under gpurun_out/r02p_*.
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
NG=$(nvidia-smi -L | wc -l)
echo "GPUs: $NG"
timeout 900 $TR --nproc-per-node $NG --master-port 29511 tools/dist_check.py > gpurun_out/r02p_dist_check$NG.log 2>&1
echo "dist_check rc=$?"; grep "DIST_CHECK\|MISMATCH" gpurun_out/r02p_dist_check$NG.log | head -3 | cut -c1-400
for sp in 1 0; do
  GSPLAT_B200_EXCHANGE_SPLIT=$sp timeout 300 $TR --nproc-per-node $NG --master-port 2952$sp bench.py --gpus $NG --steps 50 --warmup 5 \
      > gpurun_out/r02p_bench${NG}_split$sp.json 2> gpurun_out/r02p_bench${NG}_split$sp.err
  echo "bench N=$NG split=$sp rc=$?"
done
for n in 4 2; do
  [ $n -le $NG ] || continue
  timeout 300 $TR --nproc-per-node $n --master-port 2953$n bench.py --gpus $n --steps 50 --warmup 5 \
      > gpurun_out/r02p_bench${n}.json 2> gpurun_out/r02p_bench${n}.err
  echo "bench N=$n rc=$?"
done
timeout 300 python bench.py --gpus 1 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r02p_bench1.json 2> gpurun_out/r02p_bench1.err
echo "bench N=1 rc=$?"
for packed in 0 1; do
  for n in $NG 4 2 1; do
    [ $n -le $NG ] || continue
    if [ $n -eq 1 ]; then
      CFG4_PACKED=$packed CFG4_STEPS=10 timeout 300 python tools/bench_cfg4.py > gpurun_out/r02p_cfg4_w${n}_packed$packed.log 2>&1
    else
      CFG4_PACKED=$packed CFG4_STEPS=10 timeout 300 $TR --nproc-per-node $n --master-port 2954$n tools/bench_cfg4.py \
          > gpurun_out/r02p_cfg4_w${n}_packed$packed.log 2>&1
    fi
    echo "cfg4 world=$n packed=$packed rc=$?"
  done
done
timeout 200 $TR --nproc-per-node $NG --master-port 29560 tools/h2d_bandwidth.py 2>/dev/null | grep H2D > gpurun_out/r02p_h2d$NG.json
grep -h CFG4 gpurun_out/r02p_cfg4_w*.log | cut -c1-300
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r02p_bench*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d.get('parity',{}).get('img_max_err'), d.get('parity',{}).get('grad_max_rel'))
    except Exception as e: print(f, 'ERR', e)
PY
cat gpurun_out/r02p_h2d$NG.json | cut -c1-600
