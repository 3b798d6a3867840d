#!/bin/bash
# scatter iteration: parity tests, bench (auto path + two-kernel pruned for comparison), launch list
set -u
mkdir -p gpurun_out
O=gpurun_out
timeout 1200 python -m pytest tests/test_gpu_scatter.py -m gpu -x -q > $O/pytest_scatter.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_scatter.log
python bench.py > $O/bench_scatter.json 2> $O/bench_scatter.err; echo "bench scatter rc=$?"; cat $O/bench_scatter.json | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['config']['path'], d['roofline']['frac'], d['roofline']['step_split_ms'], d['e2e']['value'], d['parity_vs_cpu_sample'])"
tail -3 $O/bench_scatter.err
for xc in 2 8; do SBX_FUSED_XC=$xc python bench.py --cpu-fields 2 --e2e-steps 1 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('xc=$xc', d['value'], d['ms_per_step'], d['roofline']['frac'])"; done
python bench.py --path pruned --cpu-fields 2 --e2e-steps 1 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('two-kernel', d['value'], d['ms_per_step'], d['roofline']['frac'])"
P="python bench.py --steps 2 --warmup 3 --e2e-steps 1 --cpu-fields 2"
$P > $O/plain_scatter.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"sbx|pruned|shell_|peaks" -c 100 --csv --log-file $O/launches_scatter.csv $P > $O/ncu_scatter.log 2>&1
echo "ncu launches rc=$?"
$P > $O/plain_scatter2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:pruned_fused -s 4 -c 1 -f -o $O/prof_fused $P > $O/ncu_fused.log 2>&1
echo "ncu full rc=$?"
