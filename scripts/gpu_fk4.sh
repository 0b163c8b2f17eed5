python -m pytest tests -x -q -m gpu 2>&1 | tail -4
python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e_from_poses']['value'], d['gpu_launches'], d['stages'])"
python scripts/kernel_roofline.py 2>/dev/null | grep '"fk_skin"' | cut -c1-200
