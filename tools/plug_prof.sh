python tools/plugin_verify.py 5 2>/dev/null | tail -1
ncu --metrics gpu__time_duration.sum --clock-control none -c 30000 --csv --log-file gpurun_out/r2_launches_plugin.csv python tools/plugin_verify.py 2 > gpurun_out/ncu_plug.log 2>&1
tail -2 gpurun_out/ncu_plug.log
