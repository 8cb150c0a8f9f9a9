python tools/train_probe.py | tail -1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/s12_train_launches.csv python tools/train_probe.py > /dev/null 2>&1
