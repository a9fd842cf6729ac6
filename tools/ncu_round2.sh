# round-2 ncu captures (run under gpurun on ONE GPU); every command first runs without ncu
set -x
python tools/verify_once.py 0 > gpurun_out/vo.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2_launches_verify.csv python tools/verify_once.py 0 > gpurun_out/ncu_vo.log 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/b2.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r2_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_b2.log 2>&1
python tools/kbench.py 5 > gpurun_out/kb.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:mmq_tc -c 1 -f -o gpurun_out/r2_mmq_tc python tools/kbench.py 5 > gpurun_out/ncu_kb.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fattn_tc_kernel -c 1 -f -o gpurun_out/r2_fattn_tc python tools/verify_once.py 0 > gpurun_out/ncu_fa.log 2>&1
ls -la gpurun_out/
