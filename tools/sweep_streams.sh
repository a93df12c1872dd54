# ms/step of the OC-SORT engine against the number of concurrent streams (latency- or throughput-bound?)
p() { python -c "import sys,json; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', j['config']['streams_per_gpu'], 'groups', j['config']['stream_groups'], round(j['ms_per_step'],4), round(j['value']/1e6,1), [round(k['avg_ms'],4) for k in j['roofline']['per_kernel']])"; }
for s in 1 148 296 592 1024 1184 2048 4096; do timeout 200 python bench.py --streams $s --groups ${G:-1} --steps 60 --warmup 10 --no-e2e --no-cpu-baseline 2>/dev/null | p ocsort; done
