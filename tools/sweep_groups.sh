p() { python -c "import sys,json; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', j['config']['stream_groups'], round(j['ms_per_step'],4), round(j['value']/1e6,1))"; }
timeout 300 python -m pytest tests/test_gpu_fullsize.py -x -q -k "stream_groups" 2>&1 | tail -5
for g in 1 2 4 8; do timeout 200 python bench.py --tracker boosttrack --groups $g --steps 30 --warmup 5 --no-e2e --no-cpu-baseline 2>/dev/null | p boosttrack; done
for g in 1 2 4 8; do timeout 300 python bench.py --tracker deepocsort --groups $g --objects 480 --det-cap 512 --track-cap 768 --emb-dim 256 --steps 10 --warmup 4 --no-e2e --no-cpu-baseline 2>/dev/null | p crowd_deepocsort; done
for g in 1 2 4 8; do timeout 300 python bench.py --tracker boosttrack --groups $g --objects 480 --det-cap 512 --track-cap 768 --emb-dim 256 --steps 10 --warmup 4 --no-e2e --no-cpu-baseline 2>/dev/null | p crowd_boosttrack; done
