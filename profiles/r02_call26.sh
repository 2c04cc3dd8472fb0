cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
D4B_HALO_TIMING=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 40 --warmup 5 --no-vcycle --no-sphere --no-punctures --no-affine --no-halo-ab --no-parity > gpurun_out/r02_halo_timing_n2.json 2> gpurun_out/r02_halo_timing_n2.err
grep "halo timing" gpurun_out/r02_halo_timing_n2.err
python -c "
import json
d=json.loads(open('gpurun_out/r02_halo_timing_n2.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'])"
