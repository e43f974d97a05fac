set -x
cd $GRAFT_REPO_ROOT
for w in ffno2d_ns256 ffno1d_ks512 fno1d_burgers1024 fno2d_ns64 fno2d_ns128 fno2d_ns256; do python bench.py --workload $w --steps 10 --warmup 3 > gpurun_out/r1f_bench_$w.json 2>/dev/null; done
python bench.py --precision tf32 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r1f_bench_ffno2d_ns256_tf32.json 2>/dev/null
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1f_bench_reference_ffno2d_ns256.json 2>/dev/null
python tools/step_kernels.py ffno2d_ns256 > gpurun_out/r1f_insitu_ffno2d_ns256.txt 2>/dev/null
SEQ=1 python tools/step_kernels.py ffno2d_ns256 > gpurun_out/r1f_insitu_seq_ffno2d_ns256.txt 2>/dev/null
python tools/step_kernels.py fno2d_ns256 > gpurun_out/r1f_insitu_fno2d_ns256.txt 2>/dev/null
python tools/step_kernels.py fno1d_burgers1024 > gpurun_out/r1f_insitu_fno1d_burgers1024.txt 2>/dev/null
RPDE_TC_DBG=9 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r1f_launches_ffno2d_ns256.csv python bench.py --steps 1 --warmup 3 --no-graph --no-e2e --no-cpu-baseline > /dev/null 2> gpurun_out/r1f_shapes_ffno2d_ns256.log
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel --launch-skip 3 -c 1 -o gpurun_out/r1f_ffgemm_3xtf32 -f python tools/tc_dbg.py 3xtf32 > gpurun_out/ncu_full.log 2>&1
python tools/tc_epi.py 3xtf32 > gpurun_out/r1f_tc_epi_3xtf32.log 2>&1
python tools/tc_epi.py tf32 > gpurun_out/r1f_tc_epi_tf32.log 2>&1
for f in gpurun_out/r1f_bench_*.json; do python - "$f" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1].split('/')[-1], round(d.get('ms_per_step',0),3), round(d['value'],1), (d.get('e2e') or {}).get('value'), (d.get('cpu_baseline') or {}).get('value'), (d.get('roofline') or {}).get('frac'), (d.get('clocks') or {}))
PY
done
