cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
D4B_MG_DEBUG=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 --no-sphere --no-punctures --no-affine --no-halo-ab --no-parity > gpurun_out/dbg.json 2> gpurun_out/dbg.err
grep "d4b mg" gpurun_out/dbg.err | head -20
python -c "
import json
d=json.loads(open('gpurun_out/dbg.json').read().strip().splitlines()[-1]); print(d['vcycle']['seconds_per_vcycle'], d['vcycle']['cuda_graph_replays_in_timed_cycles'])"
