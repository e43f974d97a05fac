set -x
cd $GRAFT_REPO_ROOT
R=r1h
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/${R}_pytest_gpu.log 2>&1; tail -3 gpurun_out/${R}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${R}_smoke.log 2>&1; tail -2 gpurun_out/${R}_smoke.log
for w in ffno2d_ns256 ffno1d_ks512 fno1d_burgers1024 fno2d_ns64 fno2d_ns128 fno2d_ns256; do timeout 300 python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/${R}_bench_$w.json 2>/dev/null; done
timeout 300 python bench.py --precision tf32 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${R}_bench_ffno2d_ns256_tf32.json 2>/dev/null
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${R}_bench_reference_ffno2d_ns256.json 2>/dev/null
timeout 300 python tools/step_kernels.py ffno2d_ns256 > gpurun_out/${R}_insitu_ffno2d_ns256.txt 2>/dev/null
timeout 300 python tools/specops_bench.py > gpurun_out/${R}_specops_bench.jsonl 2>/dev/null
RPDE_TC_DBG=9 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${R}_launches_ffno2d_ns256.csv python bench.py --steps 1 --warmup 3 --no-graph --no-e2e --no-cpu-baseline > /dev/null 2> gpurun_out/${R}_shapes_ffno2d_ns256.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/${R}_launches_rollout_fno1d.csv python tools/rollout_bench.py fno1d_burgers1024 > /dev/null 2>&1
timeout 450 python tools/rollout_bench.py > gpurun_out/${R}_rollout_bench.jsonl 2>/dev/null
timeout 200 python tools/step_kernels.py fno1d_burgers1024 > gpurun_out/${R}_insitu_fno1d_burgers1024.txt 2>/dev/null
timeout 300 ncu --set full --clock-control none --import-source on -k regex:fno1d_ --launch-skip 12 -c 3 -o gpurun_out/${R}_fno1d_fused -f python tools/step_kernels.py fno1d_burgers1024 > gpurun_out/${R}_ncu_fno1d_fused.log 2>&1
for f in gpurun_out/${R}_bench_*.json; do python - "$f" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1].split('/')[-1], round(d.get('ms_per_step',0),3), round(d['value'],1), (d.get('e2e') or {}).get('value'), (d.get('cpu_baseline') or {}).get('value'), (d.get('roofline') or {}).get('frac'), (d.get('clocks') or {}))
PY
done
