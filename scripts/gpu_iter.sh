timeout 600 python -m pytest tests/test_gpu_model.py -m gpu -x -q 2>&1 | tail -2
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_iter.json 2>gpurun_out/bench_iter.err; tail -3 gpurun_out/bench_iter.err; python -c "
import json; d=json.load(open('gpurun_out/bench_iter.json')); print(d['value'], d['ms_per_step'], d['stages'], d['roofline']['achieved'], 'e2e', d['e2e']['value'])"
