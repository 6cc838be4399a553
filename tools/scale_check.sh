N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 --no-cpu --no-extras --no-push > gpurun_out/r2_scale_g${N}_s20.json 2> gpurun_out/r2_scale_g${N}_s20.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 200 --warmup 10 --no-cpu --no-extras --no-push > gpurun_out/r2_scale_g${N}_s200.json 2> gpurun_out/r2_scale_g${N}_s200.err
python bench.py --steps 20 --warmup 3 --no-cpu --no-extras --no-push > gpurun_out/r2_scale_g1_s20.json 2>/dev/null
for f in gpurun_out/r2_scale_g${N}_s20.json gpurun_out/r2_scale_g${N}_s200.json gpurun_out/r2_scale_g1_s20.json; do python -c "
import json,sys
d=json.load(open('$f')); print('$f', d['n_gpus'], 'ms/step', round(d['ms_per_step'],4), 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['config']['dp_transport'], 'kern', {k:round(v['avg_ms']*1e3,1) for k,v in d['roofline']['kernels'].items()})
"; done
