N=$1
mkdir -p gpurun_out
export GSB_BENCH_EXTRAS=1
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/r2_scale_n$N.json 2> gpurun_out/r2_scale_n$N.err; echo rc=$?
grep -E "parity" gpurun_out/r2_scale_n$N.err | head -1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/r2_ref_n$N.json 2>/dev/null
python -c "
import json;d=json.load(open('gpurun_out/r2_scale_n$N.json'));print(d['ms_per_step'],d['ms_per_step_eager'],d['value'], d['e2e']['ms_per_step'], d['nccl_parity']['ok']);e=d['exchange_ms'];print({k:v for k,v in e.items() if k!='per_rank'});c=d.get('c5_sequence',{});print(c.get('error'), c.get('first_frame'), c.get('timesteps')); print(c.get('at_2M_gaussians',{}).get('first_frame'), c.get('at_2M_gaussians',{}).get('timesteps'))
r=json.load(open('gpurun_out/r2_ref_n$N.json'));print('ref', r['value'], r['cpu_baseline']['cores'])"
