set -x
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 8 --master-port 29511 bench.py --gpus 8 --steps 8 --warmup 3 > gpurun_out/bench_r2_n8_fp32.json 2> gpurun_out/bench_r2_n8_fp32.err; tail -c 400 gpurun_out/bench_r2_n8_fp32.json
$TR --nproc-per-node 8 --master-port 29512 bench.py --gpus 8 --steps 8 --warmup 3 --payload adaptive > gpurun_out/bench_r2_n8_adaptive.json 2> gpurun_out/bench_r2_n8_adaptive.err
$TR --nproc-per-node 8 --master-port 29513 bench.py --gpus 8 --steps 6 --warmup 3 --impl eager > gpurun_out/bench_r2_n8_eager.json 2> gpurun_out/bench_r2_n8_eager.err
$TR --nproc-per-node 8 --master-port 29514 tools/bench_decode.py --reps 1 --out gpurun_out/decode_r2_n8.json > gpurun_out/decode_r2_n8.log 2>&1
CUDA_VISIBLE_DEVICES=0,1,2,3 $TR --nproc-per-node 4 --master-port 29515 bench.py --gpus 4 --steps 8 --warmup 3 > gpurun_out/bench_r2_n4_fp32.json 2> gpurun_out/bench_r2_n4_fp32.err
CUDA_VISIBLE_DEVICES=0,1,2,3 $TR --nproc-per-node 4 --master-port 29516 bench.py --gpus 4 --steps 6 --warmup 3 --impl eager > gpurun_out/bench_r2_n4_eager.json 2> gpurun_out/bench_r2_n4_eager.err
NCCL_DEBUG=INFO $TR --nproc-per-node 8 --master-port 29517 tests/multigpu/check_group.py adaptive ttt 2>&1 | grep -iE "NVLS|check|ok|fail" | head -8 > gpurun_out/group_n8_check.log
for f in gpurun_out/bench_r2_n*.json gpurun_out/decode_r2_n8.json; do echo $f; python -c "
import json,sys
for l in open('$f'):
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print(d.get('n_gpus'), d.get('impl'), round(d['value'],1), d.get('ms_per_step') or d.get('ms_per_1024_tokens'), (d.get('e2e') or {}).get('value'), d.get('exposed_averaging_ms'))
"; done
