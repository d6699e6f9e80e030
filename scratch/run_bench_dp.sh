#!/bin/bash
# headline line only (S model, batch 4096 per GPU) at N GPUs; PULL=1 adds the publish + remote reads exchange; L=1 adds config 4's share
# usage: [PULL=1] [L=1] run_bench_dp.sh N
N=${1:-2}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $T --master-port 29671 bench.py --gpus $N --skip-extras --skip-e2e --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/bench_dp${N}_ll.json
[ -n "$PULL" ] && TSB200_DP_LL=0 timeout 400 $T --master-port 29672 bench.py --gpus $N --skip-extras --skip-e2e --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/bench_dp${N}_pull.json
python - <<PY
import json
for k in ('ll', 'pull'):
    try:
        d = json.loads(open('gpurun_out/bench_dp${N}_%s.json' % k).read())
        print(k, 'value %.1f M samples/s' % (d['value'] / 1e6), 'ms_per_step %.3f' % d['ms_per_step'], d['config'].get('dp_mode'), d['config'].get('dp_mode_trial_us_per_batch'), d.get('clocks', {}).get('sm_mhz'))
    except Exception as e:
        print(k, 'failed', e)
PY
if [ -n "$L" ]; then
  timeout 300 $T --master-port 29673 scratch/ldp.py nvls $((65536 / N)) 2>&1 | grep "L model"
fi
