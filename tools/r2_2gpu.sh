#!/bin/bash
O=gpurun_out; T=${1:-r02v}
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 2 --reads 60000 --method-reads 30000 --e2e-gb 1.0 ) > $O/bench_2gpu_$T.json 2> $O/bench_2gpu_$T.err; echo "bench exit $?"; tail -5 $O/bench_2gpu_$T.err
python -c "
import json
b=json.loads(open('$O/bench_2gpu_$T.json').read().strip().split('\n')[-1])
print('value', b['value'], 'n_gpus', b['n_gpus'], 'strong', b['scaling_strong'] and {k:b['scaling_strong'][k] for k in ('value','efficiency_vs_weak_per_gpu','imbalance')})
print('e2e', b['e2e']['value'], b['e2e'].get('link_ceiling_gbs'))
print({m:(round(v['value']),round(v['roofline']['frac'],3)) for m,v in b['methods'].items()})"
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 ) > $O/bench_2gpu_ref_$T.json 2> $O/bench_2gpu_ref_$T.err; echo "ref exit $?"; head -c 300 $O/bench_2gpu_ref_$T.json
