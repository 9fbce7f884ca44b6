for n in 1 2 4 8; do
  if [ $n = 1 ]; then python bench.py --gpus 1 --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/r02_scale_n$n.json 2>/dev/null;
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 100 --warmup 10 > gpurun_out/r02_scale_n$n.json 2>gpurun_out/r02_scale_n$n.err; fi
  echo "weak n=$n rc=$?"; python -c "
import json
d=json.loads(open('gpurun_out/r02_scale_n$n.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['roofline']['frac_of_nominal_8TBs'], d['exchange_check'], d['e2e']['value'], d['api_device']['ms_per_step'])"
done
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 100 --warmup 10 --scaling strong > gpurun_out/r02_strong_n$n.json 2>/dev/null
  echo "strong n=$n rc=$?"; python -c "
import json
d=json.loads(open('gpurun_out/r02_strong_n$n.json').read().strip().splitlines()[-1])
print(d['ms_per_step'], d['value'], d['exchange_check'], d['roofline']['kernel'][:120])"
done
for n in 4 8; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n benchmarks/optim_bench.py --sizes 1600000,16000000,64000000 --out gpurun_out/r02_optim_n$n.jsonl 2>&1 | grep -E "impl" | cut -c1-230
done
python -m pytest tests/test_exchange_gpu.py tests/test_optim.py -m gpu -q 2>&1 | tail -2
