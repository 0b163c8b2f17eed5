set -x
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py tests/test_gpu_crop_cameras.py tests/test_gpu_tc.py -m gpu -q > gpurun_out/pytest_lattice.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/pytest_lattice.log
for f in 1 0; do
  AT_CROP_LATTICE=$f timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_lattice_$f.json 2> gpurun_out/bench_lattice_$f.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_lattice_$f.json"))
    print("lattice=$f", "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "poses", round(d["e2e_from_poses"]["value"]), "ms", round(d["ms_per_step"],4), "crop_ms", round(d["stages"]["crop_ms"],4), "crop_frac", round(d["stages"]["crop_hbm_frac"],4))
except Exception as e:
    print("bench failed", e); print(open("gpurun_out/bench_lattice_$f.err").read()[-1500:])
PY
done
AT_CROP_LATTICE=1 timeout 200 python scripts/kernel_roofline.py 2>&1 | grep crop_resample
