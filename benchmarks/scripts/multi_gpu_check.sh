python -m pytest tests/test_exchange_gpu.py tests/test_optim.py tests/test_distributed_cpu.py -m gpu -q 2>&1 | tail -15
echo "== bench 2 GPUs weak (lag 1)"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 100 --warmup 10 > gpurun_out/r2f_bench_n2.json 2> gpurun_out/r2f_bench_n2.err; echo rc=$?; python -c "
import json
d=json.loads(open('gpurun_out/r2f_bench_n2.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['exchange_check'], d['run']['parallelism'][:120])"
tail -3 gpurun_out/r2f_bench_n2.err
echo "== bench 2 GPUs weak (lag 0)"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 100 --warmup 10 --exchange-lag 0 > gpurun_out/r2f_bench_n2_lag0.json 2> gpurun_out/r2f_bench_n2_lag0.err; echo rc=$?; python -c "
import json
d=json.loads(open('gpurun_out/r2f_bench_n2_lag0.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['exchange_check'])"
echo "== bench 1 GPU"
python bench.py --steps 100 --warmup 10 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'])"
echo "== strong 2 GPUs"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 100 --warmup 10 --scaling strong 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['exchange_check'], d['roofline']['kernel'][:100])"
echo "== optim bench 2 GPUs"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 benchmarks/optim_bench.py --sizes 1600000,16000000 --out gpurun_out/r2f_optim_n2.jsonl 2>&1 | tail -6
