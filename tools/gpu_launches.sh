mkdir -p gpurun_out
python tools/prof_step.py vgg11 256 3 > gpurun_out/plain_p.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_cur.csv python tools/prof_step.py vgg11 256 4 > gpurun_out/ncu_l.log 2>&1; echo rc=$?
python tools/launch_summary.py gpurun_out/launches_cur.csv | head -40
