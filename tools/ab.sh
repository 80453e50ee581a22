python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -1
for i in 1 2; do python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], {k:round(v['ms'],3) for k,v in d['stages'].items()})"; done
python bench.py --forward-only --points 6000000 --width 3840 --height 2160 --steps 20 --warmup 5 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], {k:round(v,3) for k,v in d['stages_ms'].items()})"
